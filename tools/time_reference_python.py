"""Time the reference's OWN Python env.step (oracle-A: /root/reference on the restated Bullet client) in the build container.

BASELINE.md section 3 item 2: per-core env-steps/s of the reference Python, with ``np.load`` of the value table as shipped
(every step, distur_gener.py:155) and hoisted (cached).  Cannot travel to the GPU box (needs /root/reference).
Usage: python tools/time_reference_python.py [steps]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'oracle')); sys.path.insert(0, ROOT)
import ref_harness as rh  # noqa: E402
from disturbance_crazyfile_simulation_b200 import hj_tables  # noqa: E402


def run(env_id, steps, hoist):
    tables = {l: hj_tables.synthetic_table(l, 15) for l in hj_tables.LEVELS if l > 0} if 'Adversary' in env_id and 'Without' not in env_id else {}
    np.random.seed(0)
    real_load, cache = np.load, {}

    def cached_load(path, *a, **k):
        if path not in cache:
            cache[path] = real_load(path, *a, **k)
        return cache[path]

    with rh.table_dir(tables):
        env = rh.make_ref_env(env_id)
        if hoist:
            np.load = cached_load
        try:
            env.reset()
            hover = 2.0 / 1.8 - 1.0
            rng = np.random.default_rng(1)
            for _ in range(20):
                env.step((hover + 0.05 * rng.standard_normal(4)).astype(np.float32))
            t0 = time.perf_counter()
            n = 0
            for _ in range(steps):
                _, _, done, _ = env.step((hover + 0.05 * rng.standard_normal(4)).astype(np.float32))
                n += 1
                if done:
                    env.reset()
            dt = time.perf_counter() - t0
        finally:
            np.load = real_load
    return n / dt


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 600
    out = {'cpu': open('/proc/cpuinfo').read().split('model name')[1].split('\n')[0].strip(': \t'), 'cores_used': 1,
           'note': 'reference Python (envs/*.py, distur_gener.py) on oracle/fake_bullet.py instead of PyBullet: Bullet solver restated'}
    for env_id in ('DroneHoverBulletFreeEnvWithoutAdversary-v0', 'DroneHoverBulletFreeEnvWithRandomHJAdversary-v0'):
        out[env_id] = {'np_load_hoisted': run(env_id, steps, True)}
        if 'RandomHJ' in env_id:
            out[env_id]['np_load_as_shipped'] = run(env_id, max(50, steps // 6), False)
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
