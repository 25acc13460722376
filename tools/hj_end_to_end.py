"""hj_end_to_end.py - full FasTrack solves on the GPU, then the env stepped against the generated tables.

    python tools/hj_end_to_end.py OUT_DIR [levels ...]
"""
import os
import subprocess
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from disturbance_crazyfile_simulation_b200 import hj_tables, make  # noqa: E402
from disturbance_crazyfile_simulation_b200.pd_actions import pd_actions  # noqa: E402

out = sys.argv[1]
levels = [float(v) for v in sys.argv[2:]] or [1.0, 2.0]
subprocess.check_call([sys.executable, os.path.join(ROOT, 'tools', 'make_hj_tables.py'), out, '--levels'] + [str(l) for l in levels])
for level in levels:
    V = np.load(os.path.join(out, hj_tables.table_filename(level)))
    S = hj_tables.synthetic_table(level)
    env = make('DroneHoverBulletFreeEnvWithAdversary-v0', num_envs=8192, seed=3, disturbance_level=level, hj_table_dir=out)
    obs = env.reset()
    act = pd_actions(obs)
    for _ in range(1000):
        obs, rew, done, info = env.step(act)
        pd_actions(obs, act)
    st = env.episode_stats()
    print(f'level {level}: generated table finite={bool(np.isfinite(V).all())} min {V.min():.4f} max {V.max():.4f} '
          f'(synthetic stand-in: {S.min():.2f}..{S.max():.2f}); PD closed loop, 8192 envs x 1000 steps: '
          f"{st['episodes']} episodes, mean length {st['mean_length']:.1f}, mean return {st['mean_return']:.1f}", flush=True)
    env.close()
