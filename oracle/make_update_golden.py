"""Golden fixture for the PPO update (row f1): the reference's OWN loss functions and optimiser steps on a seeded batch.

TEST INFRASTRUCTURE ONLY; runs in the build container (needs /root/reference), writes tests/golden/xcheck/ppo_update.npz.
Executed, not restated: ``ActorCritic`` / ``MLPGaussianActor`` / ``MLPCritic`` (algs/core.py:176-412),
``ProximalPolicyOptimizationAlgorithm.compute_loss_pi`` (algs/ppo/ppo.py:22-40), ``IWPGAlgorithm.compute_loss_v``
(algs/iwpg/iwpg.py:269-271), ``core.get_optimizer`` (core.py:22-27: torch Adam with default betas / eps) - called on a
stand-in ``self`` that carries only the attributes those methods read (``ac``, ``clip_ratio``, ``entropy_coef``).
Recorded: the initial state_dict, the batch, loss / gradient of the policy and value losses, and the parameters after 1 and
3 optimiser steps of each network (policy: full batch, iwpg.py:448-459; value: one mini-batch per step, iwpg.py:492-505).
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import ref_harness as rh  # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden', 'xcheck', 'ppo_update.npz')


def main():
    rh.install()
    import torch
    import gym
    from phoenix_drone_simulation.algs import core
    from phoenix_drone_simulation.algs.ppo import ppo
    from phoenix_drone_simulation.algs.iwpg import iwpg
    torch.manual_seed(1234)
    np.random.seed(1234)
    od, B = 34, 512
    obs_space = gym.spaces.Box(-1000 * np.ones(od, np.float32), 1000 * np.ones(od, np.float32))
    act_space = gym.spaces.Box(-np.ones(4, np.float32), np.ones(4, np.float32))
    ac_kwargs = dict(pi=dict(activation='relu', hidden_sizes=[50, 50]), val=dict(activation='tanh', hidden_sizes=[64, 64]))
    ac = core.ActorCritic('mlp', obs_space, act_space, ac_kwargs, use_standardized_obs=True, use_scaled_rewards=True)
    ac.pi.set_log_std(0.7)
    me = types.SimpleNamespace(ac=ac, clip_ratio=0.2, entropy_coef=0.01)
    obs = torch.randn(B, od)
    with torch.no_grad():
        mu = ac.pi.net(obs)
        act = mu + ac.pi.std * torch.randn(B, 4)
        logp_old = ac.pi.log_prob_from_dist(ac.pi.dist(obs), act) + 0.15 * torch.randn(B)    # ratios on both sides of the clip
    adv = torch.randn(B) * 3.0
    target_v = torch.randn(B) * 2.0
    data = dict(obs=obs, act=act, adv=adv, log_p=logp_old, target_v=target_v)
    out = {f'init/{k}': v.detach().numpy().copy() for k, v in ac.state_dict().items()}
    out.update({f'batch/{k}': v.numpy().copy() for k, v in data.items()})
    # ActorCritic.step on raw observations in eval mode (core.py:371-394) with non-trivial running statistics
    import copy
    ac0, ac = ac, copy.deepcopy(ac)
    ac.obs_oms.mean.data = torch.randn(od) * 0.2
    ac.obs_oms.std.data = torch.rand(od) + 0.5
    ac.eval()
    raw = torch.randn(64, od) * 2
    a, v, lp = ac.step(raw)
    out['step/obs_mean'], out['step/obs_std'] = ac.obs_oms.mean.numpy().copy(), ac.obs_oms.std.numpy().copy()
    out['step/raw_obs'], out['step/act'], out['step/val'], out['step/logp'] = raw.numpy(), a, v, lp
    ac = ac0
    ac.train()
    pi_opt = core.get_optimizer('Adam', module=ac.pi, lr=3e-4)
    v_opt = core.get_optimizer('Adam', module=ac.v, lr=1e-3)
    for step in range(3):
        pi_opt.zero_grad()
        loss_pi, info = ppo.ProximalPolicyOptimizationAlgorithm.compute_loss_pi(me, data)
        loss_pi.backward()
        if step == 0:
            out['pi/loss'] = np.array(loss_pi.item())
            out['pi/ratio_mean'] = np.array(info['ratio'])
            out['pi/entropy'] = np.array(info['ent'])
            out['pi/grad'] = np.concatenate([p.grad.numpy().ravel() for p in ac.pi.net.parameters()])
        pi_opt.step()
        if step in (0, 2):
            out[f'pi/params_after_{step + 1}'] = np.concatenate([p.detach().numpy().ravel() for p in ac.pi.net.parameters()])
    mbs = B // 4
    for step in range(3):
        idx = np.arange(step * mbs, (step + 1) * mbs)
        v_opt.zero_grad()
        loss_v = iwpg.IWPGAlgorithm.compute_loss_v(me, obs=obs[idx], ret=target_v[idx])
        loss_v.backward()
        if step == 0:
            out['v/loss'] = np.array(loss_v.item())
            out['v/grad'] = np.concatenate([p.grad.numpy().ravel() for p in ac.v.parameters()])
        v_opt.step()
        if step in (0, 2):
            out[f'v/params_after_{step + 1}'] = np.concatenate([p.detach().numpy().ravel() for p in ac.v.parameters()])
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, **out)
    print('wrote', OUT, os.path.getsize(OUT) // 1024, 'KiB', 'loss_pi', out['pi/loss'], 'loss_v', out['v/loss'])


if __name__ == '__main__':
    main()
