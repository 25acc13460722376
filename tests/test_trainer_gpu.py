"""GPU: the batched PPO trainer with the fused rollout kernel and CUDA-graph replays (row f1, BASELINE.json configs[4])."""
import copy

import numpy as np
import pytest

from disturbance_crazyfile_simulation_b200.vec_env import make
from helpers import rel_err

pytestmark = pytest.mark.gpu
ENV = 'DroneHoverBulletFreeEnvWithRandomHJAdversary-v0'


def test_fused_rollout_buffers_are_what_the_torch_modules_give():
    """Rollout of T steps through the fused kernel (direct launches in epoch 0, one graph replay in epoch 1): values and
    log-probabilities in the buffers equal the torch modules evaluated on the buffered observations / actions; every row's
    next observation is the following row; time-limit cuts bootstrap from V(terminal observation), true terminals from 0."""
    import torch
    from disturbance_crazyfile_simulation_b200.trainer import PPOTrainer
    env = make(ENV, num_envs=1024, seed=5, return_terminal_obs=True, max_episode_steps=12)    # short limit: many truncations
    t = PPOTrainer(env, rollout_len=16, epochs=10, train_pi_iterations=2, train_v_iterations=1, num_mini_batches=4, seed=0)
    assert t.fused and t.use_graphs
    for epoch in range(3):
        t.roll_out()
        T, N = t.T, t.N
        ac = t.ac
        with torch.no_grad():
            o = ac.obs_oms(t.obs_buf.reshape(T * N, -1))
            v_ref = ac.v(o).reshape(T, N)
            lp_ref = ac.pi.log_prob(ac.pi.net(o), t.act_buf.reshape(T * N, 4)).reshape(T, N)
        assert rel_err(t.val_buf.cpu().numpy(), v_ref.cpu().numpy()) < 1e-5
        assert np.max(np.abs(t.logp_buf.cpu().numpy() - lp_ref.cpu().numpy())) < 2e-4
        with torch.no_grad():
            lv = ac.value(t._obs_all[T])
        assert rel_err(t.last_val.cpu().numpy(), lv.cpu().numpy()) < 1e-5
        trunc = t.trunc_buf.bool()
        assert trunc.any() and (t.done_buf.bool() & ~trunc).any()
        assert torch.all(t.end_val_buf[~trunc] == 0)
        with torch.no_grad():
            ev = ac.value(t.term_buf[trunc])
        assert rel_err(t.end_val_buf[trunc].cpu().numpy(), ev.cpu().numpy()) < 1e-5
        # exploration noise differs from step to step and from env to env
        z = ((t.act_buf - ac.pi.net(o).reshape(T, N, 4)) / torch.exp(ac.pi.log_std)).detach()
        assert abs(float(z.mean())) < 0.02 and abs(float(z.std()) - 1.0) < 0.02
        assert not torch.equal(z[0], z[1])
        t.epoch += 1                                  # epoch >= 1: the rollout is a graph replay
    assert t._rollout_graph is not None
    env.close()


def test_graph_replayed_update_equals_the_eager_update():
    import torch
    from disturbance_crazyfile_simulation_b200.trainer import PPOTrainer
    env = make(ENV, num_envs=1024, seed=5, return_terminal_obs=True)
    kw = dict(rollout_len=16, epochs=10, train_pi_iterations=5, train_v_iterations=2, num_mini_batches=4, seed=0)
    a = PPOTrainer(env, use_cuda_graphs=True, **kw)
    b = PPOTrainer(env, use_cuda_graphs=False, fused_rollout=True, **kw)
    b.ac.load_state_dict(copy.deepcopy(a.ac.state_dict()))
    b._fp.rebind()
    a.roll_out()
    for name in ('_obs_all', 'act_buf', 'rew_buf', 'val_buf', 'logp_buf', 'end_val_buf', 'done_buf', 'last_val'):
        getattr(b, name).copy_(getattr(a, name))
    for rep in range(2):                              # second pass: the graphs are replayed, not captured
        torch.manual_seed(11 + rep); ia = a.update()
        torch.manual_seed(11 + rep); ib = b.update()
        for (k, p), q in zip(a.ac.state_dict().items(), b.ac.state_dict().values()):
            assert rel_err(p.cpu().numpy(), q.cpu().numpy()) < 2e-5, (rep, k)
        assert abs(ia['loss_pi'] - ib['loss_pi']) < 1e-4 * max(1.0, abs(ib['loss_pi']))
        assert abs(ia['loss_v_before'] - ib['loss_v_before']) < 1e-4 * max(1.0, abs(ib['loss_v_before']))
    env.close()


def test_trainer_learns_to_survive_longer():
    """A few epochs at small scale: mean episode length grows (the random policy crashes after ~30-80 steps)."""
    from disturbance_crazyfile_simulation_b200.trainer import PPOTrainer
    env = make('DroneHoverBulletFreeEnvWithoutAdversary-v0', num_envs=4096, seed=1, return_terminal_obs=True)
    t = PPOTrainer(env, rollout_len=32, epochs=40, train_pi_iterations=20, train_v_iterations=3, num_mini_batches=8, seed=2)
    first = t.learn_one_epoch()
    for _ in range(24):
        last = t.learn_one_epoch()
    print(f'EpLen {first["EpLen"]:.1f} -> {last["EpLen"]:.1f}, EpRet {first["EpRet"]:.0f} -> {last["EpRet"]:.0f}')
    assert np.isfinite(last['loss_pi']) and last['EpLen'] > 1.3 * first['EpLen']
    env.close()


def test_penalty_terms_and_episode_rows():
    """penalty_*_log of the reference (hover_free.py:229-233) recomputed on the host add up to the kernel's reward; the
    per-episode rows (return, length, level of the finished episode) agree with what the step outputs say."""
    import torch
    n = 512
    env = make(ENV, num_envs=n, seed=21)
    obs = env.reset()
    ep_ret = torch.zeros(n, device='cuda', dtype=torch.float64)
    ep_len = torch.zeros(n, dtype=torch.int64, device='cuda')
    lvl = torch.tensor(env.disturbance_level, device='cuda')
    fin = {}
    g = torch.Generator(device='cuda').manual_seed(0)
    for t in range(60):
        a = 0.111 + 0.3 * torch.randn((n, 4), device='cuda', generator=g)
        lvl_before = torch.tensor(env.disturbance_level, device='cuda') if t % 10 == 0 else lvl_before
        lvl_now = torch.tensor(env.disturbance_level, device='cuda')
        obs, rew, done, info = env.step(a)
        ep_ret += rew.double(); ep_len += 1
        for i in done.nonzero().flatten().tolist():
            fin[i] = (float(ep_ret[i]), int(ep_len[i]), float(lvl_now[i]))
        ep_ret[done] = 0; ep_len[done] = 0
        if t == 7:                                   # auto-reset envs show their NEW state: compare the ones still flying
            pen = env.penalty_terms(a.cpu().numpy(), (done & ~info['truncated']).cpu().numpy())
            alive = ~done.cpu().numpy()
            total = pen['penalty'] + pen['penalty_z']
            assert np.max(np.abs(total[alive] + rew.cpu().numpy()[alive])) < 1e-3
            assert np.all(pen['penalty_crash'][alive] == 0)
    ret, ln, lv = env.last_episodes()
    assert len(fin) > 50
    for i, (r, l, level) in fin.items():
        assert ln[i] == l and abs(ret[i] - r) < 1e-2 * max(1.0, abs(r)) and abs(lv[i] - level) < 1e-6, (i, ret[i], r, ln[i], l, lv[i], level)
    assert np.all(ln[[i for i in range(n) if i not in fin]] == 0)
    env.close()
