#!/bin/bash
mkdir -p gpurun_out
for G in 0 1; do
python bench.py --steps 20 --warmup 3 --no-cpu-baseline --config3-steps 0 --other-configs 0 --graphs $G > gpurun_out/w$G.json 2>gpurun_out/w$G.err
python - <<PY
import json
d=json.loads(open("gpurun_out/w$G.json").read().strip().splitlines()[-1])
print("graphs=$G value",round(d["value"]),"wrapper ms",d["wrapper_path"]["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "eager", d.get("eager"))
PY
done
