#!/bin/bash
# round 2, call C: sort variants + bench after the pinned-ring fix
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sort.py -m gpu -x -q > gpurun_out/r2c_pytest_sort.log 2>&1
echo "sort pytest rc=$?"; tail -3 gpurun_out/r2c_pytest_sort.log
timeout 600 python bench.py --no-cpu-baseline --wrapper-steps 0 > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2c_bench.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'])
for k,v in d['stages'].items(): print(' ',k, round(v['ms_per_launch'],4))
PY
for DEFS in "" "-DSB_RS_THREADS=512 -DSB_RS_MINB=1" "-DSB_RS_THREADS=256 -DSB_RS_IPT32=8 -DSB_RS_MINB=3" "-DSB_RS_THREADS=256 -DSB_RS_LOOKBACK=32" "-DSB_RS_THREADS=384 -DSB_RS_IPT32=12 -DSB_RS_MINB=1" "-DSB_RS_THREADS=256 -DSB_RS_IPT32=24 -DSB_RS_MINB=1"; do
  SB_NVCC_DEFS="$DEFS" python -m splatformer_b200.build --force > /dev/null 2> gpurun_out/variant_build.err || { echo "build failed: $DEFS"; tail -5 gpurun_out/variant_build.err; continue; }
  echo "== $DEFS"
  timeout 300 python scripts/bench_sort.py 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
for k,v in d.items(): print('  ',k,{a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items()})
"
done
python -m splatformer_b200.build --force > /dev/null 2>&1
