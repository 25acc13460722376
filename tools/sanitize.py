"""sanitize.py - small end-to-end run of the step / reset / GAE kernels for compute-sanitizer:

    compute-sanitizer --tool memcheck  python tools/sanitize.py
    compute-sanitizer --tool racecheck python tools/sanitize.py

Covers: partial tiles, HJ gather, in-kernel resets (short episodes from random actions), the prefetch hint,
K-step launches and GAE.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from disturbance_crazyfile_simulation_b200 import make  # noqa: E402
from disturbance_crazyfile_simulation_b200.gae import gae  # noqa: E402

env_id, n = 'DroneHoverBulletFreeEnvWithRandomHJAdversary-v0', 1000
a = make(env_id, num_envs=n, seed=1, return_terminal_obs=True, hj_tables='synthetic')
b = make(env_id, num_envs=n, seed=1, env_id_base=n, share_tables_with=a)
acts = [torch.rand((n, 4), device='cuda') * 2 - 1 for _ in range(2)]
a.set_prefetch_next(b, acts[1]); b.set_prefetch_next(a, acts[0])
a.reset(); b.reset()
dones = 0
rews, ds = [], []
for t in range(40):
    for e, act in ((a, acts[0]), (b, acts[1])):
        o, r, d, info = e.step(act)
        dones += int(d.sum())
    rews.append(r.clone()); ds.append(d.to(torch.uint8).clone())
actk = torch.rand((4, n, 4), device='cuda') * 0.2
a.set_prefetch_next(None)
a.step_k(actk)
rew, done = torch.stack(rews), torch.stack(ds)
gae(rew, torch.zeros_like(rew), done, None, torch.zeros(n, device='cuda'))
torch.cuda.synchronize()
print(f'sanitize run ok: {dones} in-kernel resets')
a.close(); b.close()
