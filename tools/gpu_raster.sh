#!/bin/bash
# raster tests + timing on one B200
out=gpurun_out/${1:-ras1}; mkdir -p $out
timeout 900 python -m pytest tests/test_raster.py -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $out/pytest.log
timeout 300 python tools/raster_bench.py > $out/raster_bench.json 2> $out/raster_bench.err; echo "bench rc=$?"; cat $out/raster_bench.json; tail -3 $out/raster_bench.err
