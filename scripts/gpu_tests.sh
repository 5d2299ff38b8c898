#!/bin/bash
# Runs on the GPU box (via gpurun): the parity suite, output kept under gpurun_out/.
mkdir -p gpurun_out
timeout ${1:-600} python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -60 gpurun_out/pytest_gpu.log
