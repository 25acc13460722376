#!/bin/bash
# HJ solver: GPU parity tests, then timing of full-size (15^6) updates
TAG=${1:-hj}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_hj_solver.py -m gpu -x -q > $OUT/pytest_hj.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
timeout 600 python - > $OUT/timing.log 2>&1 <<'PY'
import time, numpy as np
from disturbance_crazyfile_simulation_b200 import hj_solver
s = hj_solver.HjSolver(hj_solver.fastrack_problem(1.5, 15))
s.run(2.0, 1e-3, max_updates=50)
t = time.time(); n, tf = s.run(2.0, 1e-3, max_updates=400); w = time.time() - t
ms, cnt = s.timing()
cells = 15 ** 6
print(f'15^6: {cnt} updates, avg device time {ms*1e3:.1f} us/update, wall {w/n*1e6:.1f} us/update incl. host loop, t={tf:.5f}')
print(f'algorithmic bytes/update = {2*4*cells/1e6:.1f} MB (read V, write V\') -> {2*4*cells/(ms*1e-3)/1e9:.0f} GB/s')
PY
echo "timing rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt; tail -15 $OUT/pytest_hj.log; cat $OUT/timing.log
