N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --wrapper-steps 0 > gpurun_out/r2r_bench_g$N.json 2> gpurun_out/r2r_bench_g$N.err; echo rc=$?; grep -v "OMP_NUM\|^\*\*\*\|^$" gpurun_out/r2r_bench_g$N.err | tail -5
python - <<PY
import json
d=json.loads(open("gpurun_out/r2r_bench_g$N.json").read().strip().splitlines()[-1])
print("N=$N value",round(d["value"]),round(d["ms_per_step"],3))
c=d["config3"]; print(" scene",{k:v for k,v in c["scene_sharded"].items() if k!="collective"}); print(" view",{k:(round(v,4) if isinstance(v,float) else v) for k,v in c["view_sharded"].items() if k!="collective"})
PY
