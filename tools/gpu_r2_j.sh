#!/bin/bash
TAG=${1:-r2j}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1200 python -m pytest tests/test_trainer_gpu.py tests/test_raster.py tests/test_policy_kernel.py tests/test_gpu_parity.py -m gpu -x -q -s > $OUT/pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
timeout 300 python tools/host_bw_probe.py > $OUT/host_bw_probe.json 2>&1
cat $OUT/rc.txt; grep -i "passed\|failed\|error\|EpLen" $OUT/pytest.log | tail -8 | cut -c1-300; cat $OUT/host_bw_probe.json | cut -c1-1500
