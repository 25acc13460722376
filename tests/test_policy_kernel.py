"""GPU: the fused rollout networks (cf_policy_step_dev, row f1) against the torch modules on the same weights.

Reference: ActorCritic.step (algs/core.py:371-394): obs_oms -> v(obs), pi.sample / pi.predict.  Tolerance 1e-5 relative
(float32 FFMA on both sides, different summation order); the exploration noise is integer work (Philox) and is compared
with oracle-B's generator."""
import ctypes as C

import numpy as np
import pytest

import cf_oracle as co
from disturbance_crazyfile_simulation_b200 import abi
from disturbance_crazyfile_simulation_b200.vec_env import make
from helpers import rel_err

pytestmark = pytest.mark.gpu


def _random_ac(od, seed, device):
    import torch
    from disturbance_crazyfile_simulation_b200.trainer import ActorCritic
    torch.manual_seed(seed)
    ac = ActorCritic(od, 4).to(device)
    with torch.no_grad():
        for p in list(ac.pi.net.parameters()) + list(ac.v.net.parameters()):
            p.mul_(1.5).add_(0.05 * torch.randn_like(p))
        ac.obs_oms.mean.data.copy_(torch.randn(od, device=device) * 0.3)
        ac.obs_oms.std.data.copy_(torch.rand(od, device=device) * 2 + 0.05)
        ac.pi.log_std.data.copy_(torch.tensor([-0.7, -1.2, 0.1, -2.5], device=device))
    return ac


@pytest.mark.parametrize('env_id,kw,n', [('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', {}, 1000),
                                          ('DroneHoverBulletFreeEnvWithoutAdversary-v0', dict(observation_noise=-1), 333)])
def test_fused_policy_matches_torch_modules(env_id, kw, n):
    import torch
    from disturbance_crazyfile_simulation_b200.policy import FusedPolicy
    torch.backends.cuda.matmul.allow_tf32 = False
    env = make(env_id, num_envs=n, seed=3, env_id_base=(5 << 32) - 400, **kw)      # ids straddle 2^32
    od = env.obs_dim
    ac = _random_ac(od, 1, env.device)
    fp = FusedPolicy(env, ac)
    obs = env.reset().clone()
    obs[::7] *= 30.0                                  # some rows far outside the standardised range (tanh saturation)
    with torch.no_grad():
        o = ac.obs_oms(obs)
        v_ref, mu_ref = ac.v(o), ac.pi.net(o)
    act, val, logp = fp.step(obs, deterministic=True)
    assert rel_err(val.cpu().numpy(), v_ref.cpu().numpy()) < 1e-5
    assert rel_err(act.cpu().numpy(), mu_ref.cpu().numpy()) < 1e-5
    assert torch.all(logp == 1.0)                     # predict(): log_p = ones (core.py:284-290)
    # sampling: a = mu + std * z with z the env's Philox event (global id, event counter, block 0, tag 4)
    act, val, logp = fp.step(obs, deterministic=False)
    std = torch.exp(ac.pi.log_std)
    z = ((act - mu_ref) / std).cpu().numpy().astype(np.float64)
    ctr = co.state_record_from_f32(env.get_state())[:, abi.STATE_FIELDS['rng_ctr'][0]].astype(np.uint64)
    seed, base = 3, (5 << 32) - 400
    z_ref = np.zeros((n, 4))
    for i in range(n):
        gid = base + i
        key = (C.c_uint32 * 2)(seed & 0xffffffff, seed >> 32)
        cw = (C.c_uint32 * 4)(gid & 0xffffffff, int(ctr[i]), 0, (4 ^ ((gid >> 32) << 8)) & 0xffffffff)
        out = (C.c_uint32 * 4)()
        co.lib().cfo_philox(cw, key, out)
        w = np.array(out[:], dtype=np.uint64)
        u = ((w >> np.uint64(9)).astype(np.float64) * 2 + 1) / 16777216.0
        r0, r1 = np.sqrt(-2 * np.log(u[0])), np.sqrt(-2 * np.log(u[2]))
        t0, t1 = np.pi * (2 * u[1] - 1), np.pi * (2 * u[3] - 1)
        z_ref[i] = [r0 * np.cos(t0), r0 * np.sin(t0), r1 * np.cos(t1), r1 * np.sin(t1)]
    # (act - mu) / std amplifies the float32 rounding of act by 1 / std (std down to e^-2.5): compare in action units
    # measured 7.9e-6: the kernel's Box-Muller (2e-6, tests/test_rng_bits.py) plus the float32 rounding of mu and act (|mu| ~ 10)
    assert np.max(np.abs((z - z_ref) * std.cpu().numpy())) < 3e-5
    lp_ref = torch.distributions.Normal(mu_ref, std).log_prob(act).sum(-1)
    assert np.max(np.abs(logp.cpu().numpy() - lp_ref.cpu().numpy())) < 1e-4 * max(1.0, float(lp_ref.abs().max()))
    # a second call without an env step in between repeats the same noise (keyed by the event counter); after a step it differs
    act2, _, _ = fp.step(obs, deterministic=False)
    assert torch.equal(act, act2)
    env.step(act)
    act3, _, _ = fp.step(obs, deterministic=False)
    assert not torch.equal(act, act3)
    env.close()


def test_reference_checkpoint_gives_the_same_actions_through_both_paths():
    import torch
    import closed_loop as cl
    from disturbance_crazyfile_simulation_b200.policy import FusedPolicy
    env = make('DroneHoverBulletFreeEnvWithAdversary-v0', num_envs=4096, seed=9)
    ac = cl.load_actor_critic('free_adv15_15431', 200, env.device)
    fp = FusedPolicy(env, ac)
    obs = env.reset()
    for _ in range(20):
        a_ref, v_ref, _ = ac.step(obs, deterministic=True)
        a, v, _ = fp.step(obs, deterministic=True)
        assert rel_err(a.cpu().numpy(), a_ref.cpu().numpy()) < 1e-5 and rel_err(v.cpu().numpy(), v_ref.cpu().numpy()) < 1e-5
        obs, _, _, _ = env.step(a)
    env.close()
