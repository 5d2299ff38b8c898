#!/bin/bash
# One GPU-box call: parity suite, smoke, bench, ncu launch list (+ optional full capture of one kernel).
#   scripts/gpu_round.sh [tag] [ncu-kernel-regex]
TAG=${1:-r1}
KREGEX=${2:-}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -5 gpurun_out/${TAG}_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/${TAG}_bench.json; tail -5 gpurun_out/${TAG}_bench.err
BENCH_SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --wrapper-steps 0"
timeout 300 $BENCH_SHORT > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $BENCH_SHORT > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu launches rc=$?"
if [ -n "$KREGEX" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KREGEX -s 4 -c 2 \
      -o gpurun_out/${TAG}_prof -f $BENCH_SHORT > gpurun_out/${TAG}_ncu_full.log 2>&1
  echo "ncu full rc=$?"
fi
