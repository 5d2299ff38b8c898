#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sort.py -m gpu -x -q > gpurun_out/r2d_pytest_sort.log 2>&1
echo "sort pytest rc=$?"; tail -3 gpurun_out/r2d_pytest_sort.log
for DEFS in "" "-DSB_RS_THREADS=512 -DSB_RS_MINB=1" "-DSB_RS_THREADS=256 -DSB_RS_IPT32=8 -DSB_RS_IPT64=8 -DSB_RS_MINB=3" "-DSB_RS_THREADS=256 -DSB_RS_IPT32=12 -DSB_RS_IPT64=8 -DSB_RS_MINB=3" "-DSB_RS_THREADS=512 -DSB_RS_IPT32=8 -DSB_RS_IPT64=6 -DSB_RS_MINB=2"; do
  SB_NVCC_DEFS="$DEFS" python -m splatformer_b200.build --force > /dev/null 2> gpurun_out/variant_build.err || { echo "build failed: $DEFS"; tail -5 gpurun_out/variant_build.err; continue; }
  echo "== $DEFS"
  timeout 300 python scripts/bench_sort.py 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
for k,v in d.items(): print('  ',k,{a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items()})
"
done
python -m splatformer_b200.build --force > /dev/null 2>&1
SORT_BENCH_ONLY=config4_47M ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r2d_sort_launches.csv python scripts/bench_sort.py > gpurun_out/r2d_sort_ncu.log 2>&1; python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/r2d_sort_launches.csv")) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value")
from collections import defaultdict
agg=defaultdict(list)
for r in rows[1:]:
    try: agg[r[ki][:70]].append(float(r[vi].replace(",","")))
    except: pass
for k,v in agg.items():
    if "sb::" in k or "cub::" in k: print(len(v), round(sum(v)/len(v)/1000,1),"us", k)
PY
