#!/bin/bash
# full parity suite + bench
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.jsonl
timeout 1500 python -m pytest tests -m gpu -q -rP --durations=8 > gpurun_out/r2f_pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2f_pytest_gpu.log
grep -E "passed|failed|error|rc=|^FAILED" gpurun_out/r2f_pytest_gpu.log | tail -8
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2f_bench.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],'wrapper',d.get('wrapper_path',{}).get('ms_per_step'))
for k,v in d['stages'].items(): print(' ',k, round(v['ms_per_launch'],4))
PY
tail -3 gpurun_out/r2f_bench.err
python scripts/host_overhead.py 2>&1 | tail -15
