#!/bin/bash
# Build-and-bench tuning variants on the GPU box.  scripts/gpu_variants.sh "<defs1>" "<defs2>" ...
mkdir -p gpurun_out
for DEFS in "$@"; do
  SB_NVCC_DEFS="$DEFS" python -m splatformer_b200.build --force > /dev/null 2> gpurun_out/variant_build.err || { echo "build failed: $DEFS"; tail -5 gpurun_out/variant_build.err; continue; }
  python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --wrapper-steps 0 --config3-steps 0 --other-configs 0 --graphs 0 > gpurun_out/variant.json 2>/dev/null
  python - "$DEFS" <<PY
import json, sys
d = json.loads(open("gpurun_out/variant.json").read().strip().splitlines()[-1])
print(f"{sys.argv[1]!r:40s} step {d['ms_per_step']:.4f} ms  " + "  ".join(f"{k}={v['ms_per_launch']:.4f}" for k, v in d["stages"].items() if v['ms_per_launch'] > 0.05))
PY
done
python -m splatformer_b200.build --force > /dev/null 2>&1
