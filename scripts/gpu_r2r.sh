#!/bin/bash
# round 2: the driver's multi-GPU launch of bench.py exactly as it will be run at round end (default flags), both arms
N=$1
mkdir -p gpurun_out
SECONDS=0
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/r_ref_g$N.json 2> gpurun_out/r_ref_g$N.err
echo "ref rc=$? seconds=$SECONDS"; tail -c 300 gpurun_out/r_ref_g$N.json
SECONDS=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r_bench_g$N.json 2> gpurun_out/r_bench_g$N.err
echo "bench rc=$? seconds=$SECONDS"; grep -v "OMP_NUM\|^\*\*\*\|^$" gpurun_out/r_bench_g$N.err | tail -5
python - <<PY
import json
d=json.loads(open("gpurun_out/r_bench_g$N.json").read().strip().splitlines()[-1])
print("N=$N value",round(d["value"]),round(d["ms_per_step"],3),"e2e",round(d["e2e"]["value"]))
c=d["config3"]; print(" scene",{k:v for k,v in c["scene_sharded"].items() if k!="collective"}); print(" view",{k:(round(v,4) if isinstance(v,float) else v) for k,v in c["view_sharded"].items() if k!="collective"})
PY
