#!/bin/bash
# NOTE: CF_HOST_SPLIT / CF_HOST_UPLOAD existed only in the experiment build this script measured (profiles/r2_e2e_timeline.md, last
# experiment); the switches are not in the tree - kept as the record of what was run.
# e2e: {single stream, split download stream} x {DMA upload, kernel upload} at depth 3 and 8, one box, two rounds
mkdir -p gpurun_out; : > gpurun_out/e2e_split.log
for rep in 1 2; do
for split in 0 1; do
for up in dma kernel; do
for d in 3 8; do
  CF_HOST_SPLIT=$split CF_HOST_UPLOAD=$up timeout 200 python bench.py --no-cpu-baseline --steps 20 --warmup 3 --e2e-depth $d > gpurun_out/e2e_s.json 2> gpurun_out/e2e_s.err
  python -c "
import json
d=json.loads([l for l in open('gpurun_out/e2e_s.json') if l.startswith('{')][-1])
print('rep $rep split $split upload $up depth $d', 'e2e %.4e'%d['e2e']['value'], 'ms/step %.4f'%(65536/d['e2e']['value']*1e3))" >> gpurun_out/e2e_split.log 2>&1 || echo "rep $rep split $split upload $up depth $d FAILED" >> gpurun_out/e2e_split.log
done; done; done; done
cat gpurun_out/e2e_split.log
CF_HOST_SPLIT=1 CF_HOST_UPLOAD=kernel timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "host" 2>&1 | tail -2
