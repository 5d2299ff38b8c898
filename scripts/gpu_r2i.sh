#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_sort.py > gpurun_out/r2i_pytest_gpu.log 2>&1
echo "pytest rc=$?"; grep -E "passed|failed|error|^FAILED" gpurun_out/r2i_pytest_gpu.log | tail -5
bash scripts/gpu_variants.sh "" "-DSB_PROJ_BWD_MINB=3" "-DSB_PROJ_FWD_MINB=4" "-DSB_PROJ_FWD_MINB=3"
