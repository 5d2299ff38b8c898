#!/bin/bash
# Round-2 profile artefacts: plain run, launch list, one --set full capture per kernel of the fused step.
mkdir -p gpurun_out
BENCH_SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --wrapper-steps 0 --config3-steps 0 --other-configs 0"
timeout 300 $BENCH_SHORT > gpurun_out/r2b_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r2b_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/r2b_final_launches.csv $BENCH_SHORT > gpurun_out/r2b_ncu_list.log 2>&1
echo "launch list rc=$?"
# warm-up 3 steps + ... : skip the first 4 steps' launches of each kernel, capture 1
timeout 1200 ncu --set full --clock-control none --import-source on \
    -k "regex:raster_rect_fwd|raster_rect_bwd|views_project_fwd|views_project_bwd|sort_tiles" -s 20 -c 5 \
    -o gpurun_out/r2b_final_prof -f $BENCH_SHORT > gpurun_out/r2b_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/r2b_ncu_full.log
