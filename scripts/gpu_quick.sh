#!/bin/bash
# tests + bench only.  scripts/gpu_quick.sh <tag> [pytest-args]
TAG=${1:-q}; shift
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q "$@" > gpurun_out/${TAG}_pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -15 gpurun_out/${TAG}_pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
print("value", round(d["value"],1), "Mpix/s  ms/step", round(d["ms_per_step"],3), " e2e", round(d["e2e"]["value"],1), "ms", round(d["e2e"]["ms_per_step"],3), "hbm_frac_step", round(d["hbm_frac_step"]["frac"],4))
for k,v in d["stages"].items(): print(f"  {k:16s} {v['ms_per_launch']:.4f} ms  share {v['share_of_step']:.3f}  frac {v.get('frac',0):.3f}")
print(d["clocks"])
PY
tail -5 gpurun_out/${TAG}_bench.err
