"""CPU: the trainer's PPO losses and optimiser steps against the reference's own (row f1).

tests/golden/xcheck/ppo_update.npz was written by oracle/make_update_golden.py, which EXECUTES the reference's
``compute_loss_pi`` (algs/ppo/ppo.py:22-40), ``compute_loss_v`` (algs/iwpg/iwpg.py:269-271), its ``ActorCritic`` and its Adam
construction (core.py:22-27) on a seeded batch.  Here the same initial ``state_dict`` is loaded into ``trainer.ActorCritic``
(strict: the key sets must be identical) and the trainer's own code must reproduce loss, gradient and the parameters after
1 and 3 Adam steps.  float32 on both sides, same torch: tolerance 1e-6 relative."""
import os
import types

import numpy as np
import pytest
import torch

from disturbance_crazyfile_simulation_b200 import trainer as tr

FIX = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'xcheck', 'ppo_update.npz')


@pytest.fixture(scope='module')
def fix():
    return dict(np.load(FIX))


def _ac(fix):
    ac = tr.ActorCritic(34, 4, use_scaled_rewards=True)
    sd = {k[len('init/'):]: torch.tensor(v) for k, v in fix.items() if k.startswith('init/')}
    ac.load_state_dict(sd, strict=True)
    return ac


def _close(a, b, tol=1e-6):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0))) < tol


def test_policy_loss_gradient_and_adam_steps(fix):
    ac = _ac(fix)
    me = types.SimpleNamespace(ac=ac, clip_ratio=0.2, entropy_coef=0.01)
    b = {k[len('batch/'):]: torch.tensor(v) for k, v in fix.items() if k.startswith('batch/')}
    opt = torch.optim.Adam(ac.pi.net.parameters(), lr=3e-4)
    for step in range(3):
        opt.zero_grad()
        loss, mu, ratio = tr.PPOTrainer.compute_loss_pi(me, b['obs'], b['act'], b['adv'], b['log_p'])
        loss.backward()
        if step == 0:
            assert _close(loss.item(), fix['pi/loss']) and _close(ratio.mean().item(), fix['pi/ratio_mean'])
            assert _close(ac.pi.entropy().item(), fix['pi/entropy'])
            assert _close(np.concatenate([p.grad.numpy().ravel() for p in ac.pi.net.parameters()]), fix['pi/grad'], 1e-6)
        opt.step()
        if step in (0, 2):
            got = np.concatenate([p.detach().numpy().ravel() for p in ac.pi.net.parameters()])
            assert _close(got, fix[f'pi/params_after_{step + 1}']), f'policy parameters after {step + 1} Adam steps'


def test_value_loss_gradient_and_adam_steps(fix):
    ac = _ac(fix)
    obs, tgt = torch.tensor(fix['batch/obs']), torch.tensor(fix['batch/target_v'])
    opt = torch.optim.Adam(ac.v.parameters(), lr=1e-3)
    mbs = obs.shape[0] // 4
    for step in range(3):
        idx = torch.arange(step * mbs, (step + 1) * mbs)
        opt.zero_grad()
        loss = ((ac.v(obs[idx]) - tgt[idx]) ** 2).mean()             # the expression of PPOTrainer.update / _v_minibatches
        loss.backward()
        if step == 0:
            assert _close(loss.item(), fix['v/loss'])
            assert _close(np.concatenate([p.grad.numpy().ravel() for p in ac.v.parameters()]), fix['v/grad'])
        opt.step()
        if step in (0, 2):
            got = np.concatenate([p.detach().numpy().ravel() for p in ac.v.parameters()])
            assert _close(got, fix[f'v/params_after_{step + 1}'])


def test_actor_critic_step_eval_mode(fix):
    ac = _ac(fix)
    ac.obs_oms.mean.data.copy_(torch.tensor(fix['step/obs_mean']))
    ac.obs_oms.std.data.copy_(torch.tensor(fix['step/obs_std']))
    a, v, _ = ac.step(torch.tensor(fix['step/raw_obs']), deterministic=True)
    assert _close(a.numpy(), fix['step/act']) and _close(v.numpy(), fix['step/val'])


def test_capturable_adam_matches_plain_adam():
    """The graph-captured update uses Adam(capturable=True, lr=tensor): on the CPU that flag is not accepted, so the
    equivalence is checked on the GPU in tests/test_trainer_gpu.py; here: the KL statistic is the all-element mean
    (iwpg.py:461-466) and the entropy the per-dimension mean (iwpg.py:259)."""
    ac = tr.ActorCritic(34, 4)
    ac.pi.log_std.data.copy_(torch.tensor([-0.5, -1.0, 0.2, -2.0]))
    ref = torch.distributions.Normal(torch.zeros(4), torch.exp(ac.pi.log_std)).entropy().mean()
    assert abs(float(ac.pi.entropy()) - float(ref)) < 1e-6
