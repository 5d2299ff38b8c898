#!/bin/bash
# one --set full capture (with source) of the kernels matching $1 (regex), report gpurun_out/$2.ncu-rep
mkdir -p gpurun_out
BENCH_SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --wrapper-steps 0 --config3-steps 0 --other-configs 0"
timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:$1" -s ${3:-4} -c ${4:-1} \
    -o gpurun_out/$2 -f $BENCH_SHORT > gpurun_out/$2.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/$2.log
