#!/bin/bash
# side kernels: rasteriser (drone pass beside the background pass) and GAE (single-wave configuration)
TAG=${1:-r2m}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_raster.py tests/test_gpu_parity.py -m gpu -x -q -k "raster or gae or capture or frames or validation" > $OUT/pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
timeout 300 python tools/raster_bench.py > $OUT/raster_bench.json 2> $OUT/raster_bench.err; echo "raster bench rc=$?" >> $OUT/rc.txt
timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 --steps 400 --warmup 50 > $OUT/bench.json 2> $OUT/bench.err
cat $OUT/rc.txt; tail -3 $OUT/pytest.log | cut -c1-300; cat $OUT/raster_bench.json | cut -c1-800; python -c "
import json
d=json.loads(open('$OUT/bench.json').read().strip().split('\n')[-1]); print(json.dumps(d['side'].get('gae')), json.dumps(d['side'].get('raster'))[:600])"
