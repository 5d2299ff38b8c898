#!/bin/bash
# one --set full capture (with source) of the fused step's five kernels, for stall reasons / per-line hot spots
mkdir -p gpurun_out
BENCH_SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --wrapper-steps 0 --config3-steps 0 --other-configs 0"
timeout 300 $BENCH_SHORT > gpurun_out/s_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/s_plain.log; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on \
    -k "regex:raster_rect_fwd|raster_rect_bwd|views_project_fwd|views_project_bwd|sort_tiles" -s 20 -c 5 \
    -o gpurun_out/r2s_prof -f $BENCH_SHORT > gpurun_out/s_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/s_ncu_full.log; ls -la gpurun_out/
