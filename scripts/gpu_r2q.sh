#!/bin/bash
# round 2, final verification: the no-allocation entry point alone in a fresh process, the whole GPU suite
# (which collects it last), smoke(), the default bench line and the reference arm.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/q_noalloc_alone.log 2>&1
echo "no_alloc alone rc=$?"; tail -3 gpurun_out/q_noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/q_gpu_suite.log 2>&1
echo "gpu suite rc=$?"; tail -3 gpurun_out/q_gpu_suite.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/q_smoke.log 2>&1
echo "smoke rc=$?"; tail -2 gpurun_out/q_smoke.log
SECONDS=0
timeout 900 python bench.py > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err
echo "bench rc=$? seconds=$SECONDS"; tail -c 1500 gpurun_out/q_bench.json
SECONDS=0
timeout 600 python bench.py --impl reference > gpurun_out/q_bench_ref.json 2> gpurun_out/q_bench_ref.err
echo "bench ref rc=$? seconds=$SECONDS"; tail -c 600 gpurun_out/q_bench_ref.json
