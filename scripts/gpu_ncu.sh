#!/bin/bash
# ncu --set full capture of our kernels in one bench step (after the warm-up steps).
#   scripts/gpu_ncu.sh <tag> <kernel-regex> <skip> <count>
TAG=${1:-r1}; KREGEX=${2:-"raster|views_project|map_isect|bin_edges"}; SKIP=${3:-18}; COUNT=${4:-6}
mkdir -p gpurun_out
BENCH_SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --wrapper-steps 0"
timeout 300 $BENCH_SHORT > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:$KREGEX" -s $SKIP -c $COUNT \
    -o gpurun_out/${TAG}_prof -f $BENCH_SHORT > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/${TAG}_ncu_full.log
