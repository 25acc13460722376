#!/bin/bash
# ncu capture of the frame renderer (plain run first, then under ncu)
out=gpurun_out/${1:-rasncu}; mkdir -p $out
timeout 300 python tools/raster_bench.py > $out/raster_bench.json 2> $out/raster_bench.err || { echo "plain run failed"; tail -5 $out/raster_bench.err; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_raster|k_background|k_frame_ctx|k_tri_setup" -s 12 -c 4 -o $out/raster -f python tools/raster_bench.py > $out/ncu.log 2>&1; echo "ncu rc=$?"; tail -3 $out/ncu.log
ncu -i $out/raster.ncu-rep --page raw --csv > $out/raster_raw.csv 2>/dev/null
