#!/bin/bash
# the render tests + the headline bench only (stage times).  scripts/gpu_fast.sh <tag>
TAG=${1:-f}
TESTS=${2:-"tests/test_gpu_fused.py tests/test_gpu_fullsize.py tests/test_gpu_parity.py"}
mkdir -p gpurun_out
timeout 900 python -m pytest $TESTS -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/${TAG}_pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline --wrapper-steps 0 --config3-steps 0 --other-configs 0 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
print("graph_replay", d.get("graph_replay"))
print("value", round(d["value"],1), "Mpix/s  ms/step", round(d["ms_per_step"],4), " e2e", round(d["e2e"]["value"],1), "ms", round(d["e2e"]["ms_per_step"],4))
for k,v in d["stages"].items(): print(f"  {k:16s} {v['ms_per_launch']:.4f} ms  share {v['share_of_step']:.3f}")
print(d["clocks"])
PY
tail -3 gpurun_out/${TAG}_bench.err
