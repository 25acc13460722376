"""make_hj_tables.py - regenerate the reference's missing FasTrack value tables on the GPU.

    python tools/make_hj_tables.py OUT_DIR [--levels 0.1 0.5 ...] [--lookback 2.0] [--t-step 0.001] [--device 0]

Writes OUT_DIR/fastrack_{level}_15x15.npy (the name distur_gener.py:155 loads) for every level (default: 0.1 .. 2.0,
the levels Boltzmann() draws) with disturbance_crazyfile_simulation_b200.hj_solver (Fastrack.py:281-320 on the GPU).
Pass OUT_DIR as ``hj_table_dir`` to ``make(...)`` to step the envs against these tables instead of the synthetic ones.
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from disturbance_crazyfile_simulation_b200 import hj_solver, hj_tables  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('out_dir')
    ap.add_argument('--levels', type=float, nargs='*', default=[l for l in hj_tables.LEVELS if l > 0])
    ap.add_argument('--lookback', type=float, default=2.0)
    ap.add_argument('--t-step', type=float, default=1e-3)
    ap.add_argument('--device', type=int, default=0)
    a = ap.parse_args()
    os.makedirs(a.out_dir, exist_ok=True)
    for level in a.levels:
        t = time.time()
        s = hj_solver.HjSolver(hj_solver.fastrack_problem(level, 15), a.device)
        n, t_final = s.run(a.lookback, a.t_step)
        V = s.get_value()
        ms, _ = s.timing()
        s.close()
        path = os.path.join(a.out_dir, hj_tables.table_filename(level))
        np.save(path, V)
        print(f'{path}: {n} updates to t = {t_final:.4f}, {ms * 1e3:.0f} us/update, {time.time() - t:.1f} s, '
              f'V in [{V.min():.4f}, {V.max():.4f}]', flush=True)


if __name__ == '__main__':
    main()
