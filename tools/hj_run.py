"""hj_run.py - a few full-size updates of the HJ solver (profiling target):  python tools/hj_run.py [updates]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from disturbance_crazyfile_simulation_b200 import hj_solver  # noqa: E402

s = hj_solver.HjSolver(hj_solver.fastrack_problem(1.5, 15))
s.run(2.0, 1e-3, max_updates=int(sys.argv[1]) if len(sys.argv) > 1 else 12)
ms, n = s.timing()
print(f'{n} updates, {ms * 1e3:.1f} us/update')
