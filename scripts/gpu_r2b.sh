#!/bin/bash
# round 2, call B: sort tests first (fail fast), parity suite, bench, sort bench
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.jsonl
timeout 900 python -m pytest tests/test_gpu_sort.py -m gpu -x -q > gpurun_out/r2b_pytest_sort.log 2>&1
echo "sort pytest rc=$?"; tail -3 gpurun_out/r2b_pytest_sort.log
timeout 1500 python -m pytest tests -m gpu -q -rP --durations=10 --deselect tests/test_gpu_sort.py > gpurun_out/r2b_pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2b_pytest_gpu.log
grep -E "passed|failed|error|rc=" gpurun_out/r2b_pytest_gpu.log | tail -5
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2b_bench.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],'wrapper',d.get('wrapper_path',{}).get('ms_per_step'))
for k,v in d['stages'].items(): print(' ',k, round(v['ms_per_launch'],4))
PY
tail -3 gpurun_out/r2b_bench.err
timeout 600 python scripts/bench_sort.py > gpurun_out/r2b_sort_bench.json 2> gpurun_out/r2b_sort_bench.err; echo "sort bench rc=$?"
cat gpurun_out/r2b_sort_bench.json; tail -3 gpurun_out/r2b_sort_bench.err
