"""World-size-2 gloo tests (CPU) of the N>1 host logic: env sharding, parameter broadcast, flattened
gradient averaging, running statistics, episode statistics, the parameter consistency check and a
PPOTrainer epoch.  The env step path itself has no collective."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from disturbance_crazyfile_simulation_b200 import dist_utils as du
from disturbance_crazyfile_simulation_b200 import trainer as tr


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _run(fn, world=2):
    port = _free_port()
    mp.spawn(_entry, args=(world, port, fn), nprocs=world, join=True)


def _entry(rank, world, port, fn):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        fn(rank, world)
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions_all_envs():
    for n, w in ((65536, 8), (1000, 3), (7, 8), (1, 1)):
        seen = []
        for r in range(w):
            b, c = du.shard_range(n, r, w)
            seen += list(range(b, b + c))
        assert seen == list(range(n))
    with pytest.raises(ValueError):
        du.shard_range(8, 2, 2)


def _collectives(rank, world):
    torch.manual_seed(rank)
    lin = torch.nn.Linear(5, 3)
    du.broadcast_params(list(lin.parameters()))
    assert du.check_distributed_parameters(lin)
    # gradient averaging: every rank ends with the mean of the per-rank gradients
    x = torch.full((4, 5), float(rank + 1))
    lin(x).sum().backward()
    g_local = [p.grad.clone() for p in lin.parameters()]
    du.FlatGradAverager(list(lin.parameters()))()
    gathered = [torch.zeros_like(torch.cat([g.reshape(-1) for g in g_local])) for _ in range(world)]
    dist.all_gather(gathered, torch.cat([g.reshape(-1) for g in g_local]))
    mean = torch.stack(gathered).mean(0)
    assert torch.allclose(torch.cat([p.grad.reshape(-1) for p in lin.parameters()]), mean, atol=1e-6)
    # consistency check trips when one rank diverges
    if rank == 1:
        with torch.no_grad():
            lin.weight.add_(1.0)
    assert not du.check_distributed_parameters(lin)
    # running statistics over the ranks == one process over the concatenated batch
    g = torch.Generator().manual_seed(5)
    full = torch.randn((2, 64, 7), generator=g) * 3.0 + 1.5
    oms = tr.OnlineMeanStd(shape=(7,))
    oms.update(full[0, rank * 32:(rank + 1) * 32] if world == 2 else full[0])
    oms.update(full[1, rank * 32:(rank + 1) * 32] if world == 2 else full[1])
    mean, std, count = torch.zeros(7), torch.ones(7), 0.0   # the reference's update (online_mean_std.py:62-95), one process
    for b in range(2):
        x = full[b]
        n_B, n_AB = float(x.shape[0]), count + x.shape[0]
        delta = x.mean(0) - mean
        mean_new = mean + delta * n_B / n_AB
        M2 = count * std ** 2 + n_B * ((x - mean_new) ** 2).mean(0) + delta ** 2 * (count * n_B / n_AB)
        mean, count, std = mean_new, n_AB, torch.sqrt(M2 / n_AB)
    assert torch.allclose(oms.mean, mean, atol=1e-5) and torch.allclose(oms.mean, full.reshape(-1, 7).mean(0), atol=1e-5)
    assert torch.allclose(oms.std, std, atol=1e-5)
    assert float(oms.count) == 128.0
    # episode statistics
    red = du.reduce_episode_stats(10 * (rank + 1), 100.0 * (rank + 1), 50.0 * (rank + 1), -5.0 - rank, 7.0 + rank, 'cpu')
    assert red['episodes'] == 30 and abs(red['mean_return'] - 10.0) < 1e-12 and abs(red['mean_length'] - 5.0) < 1e-12
    assert red['min_return'] == -6.0 and red['max_return'] == 8.0


def test_collectives_world2():
    _run(_collectives)


class _FakeEnv:
    """CPU stand-in with the CrazyFlieVecEnv surface the trainer uses (sharded by global env id)."""

    def __init__(self, n, base):
        self.num_envs, self.obs_dim, self.device = n, 34, torch.device('cpu')
        self._term = torch.zeros((n, 34))
        self._gen = torch.Generator().manual_seed(100 + base)
        self._t = torch.zeros(n, dtype=torch.long)
        self._eps = [0, 0.0, 0.0]

    def _obs(self):
        return torch.randn((self.num_envs, 34), generator=self._gen)

    def reset(self):
        return self._obs()

    def step(self, a):
        self._t += 1
        rew = -a.pow(2).sum(-1)
        done = (self._t % 7 == 0) | (torch.rand(self.num_envs, generator=self._gen) < 0.05)
        trunc = (self._t % 7 == 0)
        self._eps[0] += int(done.sum()); self._eps[1] += float(rew[done].sum()); self._eps[2] += float(self._t[done].sum())
        self._t[done] = 0
        self._term = self._obs()
        return self._obs(), rew, done, {'cost': torch.zeros_like(rew), 'truncated': trunc & done, 'terminal_observation': self._term}

    def episode_stats(self, reset=False):
        out = {'episodes': self._eps[0], 'return_sum': self._eps[1], 'length_sum': self._eps[2]}
        if reset:
            self._eps = [0, 0.0, 0.0]
        return out


def _gae_cpu(rew, val, done, end_val, last_val, gamma, lam, rew_scale):
    import cf_oracle as co                                  # the CPU oracle as the stand-in for cf_gae_dev
    a, t, r = co.gae(rew.numpy(), val.numpy(), done.numpy(), None if end_val is None else end_val.numpy(),
                     last_val.numpy(), gamma, lam, rew_scale)
    return torch.from_numpy(a), torch.from_numpy(t), torch.from_numpy(r)


def _trainer(rank, world):
    base, cnt = du.shard_range(64, rank, world)
    env = _FakeEnv(cnt, base)
    import tempfile
    from disturbance_crazyfile_simulation_b200.run_logger import RunLogger
    log_dir = tempfile.mkdtemp()
    t = tr.PPOTrainer(env, rollout_len=8, epochs=4, train_pi_iterations=3, train_v_iterations=2, num_mini_batches=4,
                      seed=3, gae_fn=_gae_cpu, check_freq=1, logger=RunLogger(log_dir, rank=rank), save_freq=2)
    w0 = torch.cat([p.detach().reshape(-1).clone() for p in t.ac.pi.net.parameters()])
    for _ in range(2):
        info = t.learn_one_epoch()                          # raises if the ranks' parameters diverge
        assert np.isfinite(info['loss_pi']) and np.isfinite(info['loss_v_before'])
        assert info['TotalEnvSteps'] == t.epoch * 8 * 64
    w1 = torch.cat([p.detach().reshape(-1) for p in t.ac.pi.net.parameters()])
    assert not torch.equal(w0, w1)
    assert float(t.ac.obs_oms.count) == 2 * 8 * 64           # both ranks' samples counted
    gathered = [torch.zeros_like(w1) for _ in range(world)]
    dist.all_gather(gathered, w1)
    assert torch.equal(gathered[0], gathered[1])
    # run directory: only rank 0 writes; two epochs -> header + 2 rows, config.json, model0.pt (the reference's numbering: epochs 0, save_freq, ...)
    import os
    if rank == 0:
        rows = open(os.path.join(log_dir, 'progress.csv')).read().strip().split('\n')
        assert len(rows) == 3 and rows[0].startswith('Epoch,EpRet/Mean,EpLen/Mean,Loss/Pi,Loss/Value')
        assert rows[2].split(',')[0] == '2' and rows[0].split(',')[-3:] == ['TotalEnvSteps', 'Time', 'FPS']
        assert os.path.exists(os.path.join(log_dir, 'config.json'))
        sd = torch.load(os.path.join(log_dir, 'torch_save', 'model0.pt'))
        assert 'pi.net.0.weight' in sd and 'obs_oms.mean' in sd
    else:
        assert os.listdir(log_dir) == []


def test_trainer_epoch_world2():
    _run(_trainer)


def test_bind_to_gpu_cpus_is_best_effort():
    """No NVML / no GPU here: the placement helper reports why it did nothing and never raises."""
    import os
    before = os.sched_getaffinity(0)
    info = du.bind_to_gpu_cpus(0)
    assert isinstance(info, dict) and info['bound'] in (True, False)
    if not info['bound']:
        assert os.sched_getaffinity(0) == before
    os.sched_setaffinity(0, before)
