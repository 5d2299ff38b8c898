#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_sort.py > gpurun_out/r2p_pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|error|^FAILED|^E  " gpurun_out/r2p_pytest_gpu.log | tail -6
python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --config3-steps 0 --other-configs 0 2>/dev/null > gpurun_out/r2p_bench.json
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2p_bench.json").read().strip().splitlines()[-1]); print("value",d["value"],d["ms_per_step"],"wrapper",d["wrapper_path"])
PY
