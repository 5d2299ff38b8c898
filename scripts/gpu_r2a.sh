#!/bin/bash
# round 2, call A: parity suite + bench
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.jsonl
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2a_smi.txt 2>&1
free -g | head -2 >> gpurun_out/r2a_smi.txt; nproc >> gpurun_out/r2a_smi.txt
timeout 1500 python -m pytest tests -m gpu -q -rP --durations=15 > gpurun_out/r2a_pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2a_pytest_gpu.log
grep -E "passed|failed|error|rc=" gpurun_out/r2a_pytest_gpu.log | tail -5
timeout 600 python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/r2a_bench.json; tail -5 gpurun_out/r2a_bench.err
