#!/bin/bash
# our arm at N GPUs, launched as the driver does
N=$1
mkdir -p gpurun_out
SECONDS=0
timeout ${2:-900} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/t_bench_g$N.json 2> gpurun_out/t_bench_g$N.err
echo "bench rc=$? seconds=$SECONDS"; grep -v "OMP_NUM\|^\*\*\*\|^$" gpurun_out/t_bench_g$N.err | tail -5
python - <<PY
import json
d=json.loads(open("gpurun_out/t_bench_g$N.json").read().strip().splitlines()[-1])
print("N=$N value",round(d["value"]),round(d["ms_per_step"],3),"e2e",round(d["e2e"]["value"]),round(d["e2e"]["ms_per_step"],3),"eager",d.get("eager"), d["config"].get("launch","")[:30])
c=d["config3"]; print(" scene",{k:v for k,v in c["scene_sharded"].items() if k!="collective"}); print(" view",{k:(round(v,6) if isinstance(v,float) else v) for k,v in c["view_sharded"].items() if k not in("collective","grad_rel_l2_per_tensor")})
PY
