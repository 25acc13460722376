"""GPU: the fused PPO policy-gradient kernel (cf_ppo_pi_grad_dev, row f1) against torch autograd on the same weights.

Reference: PPO.compute_loss_pi (algs/ppo/ppo.py:22-40) + IWPG.update_policy_net's backward (algs/iwpg/iwpg.py:448-459):
loss = -mean(min(ratio*adv, clip(ratio, 1-eps, 1+eps)*adv)), ratio = exp(logp(mu(obs), act) - logp_old); the entropy term
depends on log_std only (requires_grad False), so it adds no gradient.  Both sides compute in float32 FFMA with different
summation orders: 2e-5 relative to the largest gradient entry of each tensor.  (Round 2's first run failed the 113 669-sample
case by one sample's worth of gradient: a ratio within rounding of the clip bound; see _batch.)"""
import numpy as np
import pytest

from disturbance_crazyfile_simulation_b200.vec_env import make
from test_policy_kernel import _random_ac

pytestmark = pytest.mark.gpu


def _batch(ac, od, B, seed, device, spread):
    import torch
    g = torch.Generator(device='cpu').manual_seed(seed)
    obs = torch.randn(B, od, generator=g).to(device)
    with torch.no_grad():
        mu = ac.pi.net(obs)
        std = torch.exp(ac.pi.log_std)
        act = mu + std * torch.randn(B, 4, generator=g).to(device)
        logp = ac.pi.log_prob(mu, act)
        # old log-probabilities around the new ones: ratios on both sides of the clip interval and inside it
        logp_old = logp + spread * torch.randn(B, generator=g).to(device)
        adv = torch.randn(B, generator=g).to(device)
        # the surrogate's gradient jumps where the ratio crosses a clip bound (weight 1 <-> 0): a sample whose ratio sits within
        # float32 rounding of a bound may fall on either side in the two implementations, so such samples get zero advantage
        ratio = torch.exp(logp - logp_old)
        near = ((ratio - 0.8).abs() < 1e-4) | ((ratio - 1.2).abs() < 1e-4)
        adv[near] = 0.0
    return obs.contiguous(), act.contiguous(), adv.contiguous(), logp_old.contiguous()


@pytest.mark.parametrize('env_id,B,spread', [('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', 128 * 300, 0.15),
                                              ('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', 4099, 0.3),
                                              ('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', 77, 0.0),
                                              ('DroneHoverBulletFreeEnvWithoutAdversary-v0', 1500, 0.15),
                                              ('DroneHoverBulletFreeEnvWithoutAdversary-v0', 128 * 296 * 3 + 5, 0.15)])
def test_fused_gradient_matches_autograd(env_id, B, spread):
    import torch
    from disturbance_crazyfile_simulation_b200.policy import FusedPolicy
    from disturbance_crazyfile_simulation_b200.trainer import PPOTrainer
    torch.backends.cuda.matmul.allow_tf32 = False
    kw = dict(observation_noise=-1) if 'Without' in env_id else {}
    env = make(env_id, num_envs=128, seed=1, **kw)
    od = env.obs_dim
    ac = _random_ac(od, 5, env.device)
    fp = FusedPolicy(env, ac)
    obs, act, adv, logp_old = _batch(ac, od, B, 11, env.device, spread)
    clip = 0.2
    # torch side: the trainer's own loss (entropy_coef 0 here; the term has no gradient either way)
    for p in ac.pi.net.parameters():
        p.grad = None
    mu = ac.pi.net(obs)
    logp = ac.pi.log_prob(mu, act)
    ratio = torch.exp(logp - logp_old)
    loss = -(torch.min(ratio * adv, adv * torch.clamp(ratio, 1 - clip, 1 + clip))).mean()
    loss.backward()
    ref = [p.grad.clone() for p in ac.pi.net.parameters()]
    frac_clipped = float(((ratio < 1 - clip) | (ratio > 1 + clip)).float().mean())
    if spread > 0.1:
        assert 0.05 < frac_clipped < 0.9
    for p in ac.pi.net.parameters():
        p.grad = torch.full_like(p, 7.0)              # the kernel overwrites, it does not accumulate
    l, r = fp.pi_grad(obs, act, adv, logp_old, clip)
    torch.cuda.synchronize()
    assert abs(float(l) - float(loss.detach())) < 2e-5 * max(1.0, abs(float(loss)))
    assert abs(float(r) - float(ratio.mean())) < 2e-5
    for p, g in zip(ac.pi.net.parameters(), ref):
        d = (p.grad - g).abs().max().item()
        assert d < 2e-5 * max(g.abs().max().item(), 1e-6), (tuple(p.shape), d, g.abs().max().item())
    # a second launch on the same inputs is bit-identical (fixed tile -> CTA assignment, fixed reduction order)
    first = [p.grad.clone() for p in ac.pi.net.parameters()]
    fp.pi_grad(obs, act, adv, logp_old, clip)
    torch.cuda.synchronize()
    for p, g in zip(ac.pi.net.parameters(), first):
        assert torch.equal(p.grad, g)
    env.close()


@pytest.mark.parametrize('env_id,n_all,start,count', [('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', 70000, 1000, 128 * 300 + 17),
                                                       ('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', 5000, 0, 77),
                                                       ('DroneHoverBulletFreeEnvWithoutAdversary-v0', 300000, 4096, 131072),
                                                       ('DroneHoverBulletFreeEnvWithoutAdversary-v0', 900, None, 900)])
def test_fused_value_gradient_matches_autograd(env_id, n_all, start, count):
    """cf_ppo_v_grad_dev: one mini-batch step of update_value_net (iwpg.py:478-510) on rows perm[start:start+count]."""
    import torch
    from disturbance_crazyfile_simulation_b200.policy import FusedPolicy
    torch.backends.cuda.matmul.allow_tf32 = False
    kw = dict(observation_noise=-1) if 'Without' in env_id else {}
    env = make(env_id, num_envs=128, seed=1, **kw)
    od, dev = env.obs_dim, env.device
    ac = _random_ac(od, 7, dev)
    fp = FusedPolicy(env, ac)
    g = torch.Generator(device='cpu').manual_seed(3)
    obs = torch.randn(n_all, od, generator=g).to(dev)
    tgt = (3.0 * torch.randn(n_all, generator=g)).to(dev)
    perm = None if start is None else torch.randperm(n_all, generator=g).to(dev)
    idx = slice(0, count) if perm is None else perm[start:start + count]
    for p in ac.v.net.parameters():
        p.grad = None
    loss = ((ac.v(obs[idx]) - tgt[idx]) ** 2).mean()
    loss.backward()
    ref = [p.grad.clone() for p in ac.v.net.parameters()]
    for p in ac.v.net.parameters():
        p.grad = torch.full_like(p, -3.0)
    l = fp.v_grad(obs, tgt, perm, start or 0, count)
    torch.cuda.synchronize()
    assert abs(float(l) - float(loss.detach())) < 2e-5 * max(1.0, abs(float(loss.detach())))
    for p, gr in zip(ac.v.net.parameters(), ref):
        d = (p.grad - gr).abs().max().item()
        assert d < 2e-5 * max(gr.abs().max().item(), 1e-6), (tuple(p.shape), d, gr.abs().max().item())
    env.close()


def test_trainer_update_with_and_without_the_fused_gradient_agree():
    import copy
    import torch
    from disturbance_crazyfile_simulation_b200.trainer import PPOTrainer
    from helpers import rel_err
    torch.backends.cuda.matmul.allow_tf32 = False
    env = make('DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', num_envs=1024, seed=5, return_terminal_obs=True)
    kw = dict(rollout_len=16, epochs=10, train_pi_iterations=8, train_v_iterations=2, num_mini_batches=4, seed=0)
    a = PPOTrainer(env, fused_update=True, **kw)
    b = PPOTrainer(env, fused_update=False, **kw)
    assert a.fused_update and not b.fused_update
    b.ac.load_state_dict(copy.deepcopy(a.ac.state_dict()))
    b._fp.rebind()
    a.roll_out()
    for name in ('_obs_all', 'act_buf', 'rew_buf', 'val_buf', 'logp_buf', 'end_val_buf', 'done_buf', 'last_val'):
        getattr(b, name).copy_(getattr(a, name))
    for rep in range(2):                              # second pass: the graphs are replayed, not captured
        torch.manual_seed(11 + rep); ia = a.update()
        torch.manual_seed(11 + rep); ib = b.update()
        for (k, p), q in zip(a.ac.state_dict().items(), b.ac.state_dict().values()):
            assert rel_err(p.cpu().numpy(), q.cpu().numpy()) < 5e-5, (rep, k)
        assert abs(ia['loss_pi'] - ib['loss_pi']) < 1e-4 * max(1.0, abs(ib['loss_pi']))
    env.close()
