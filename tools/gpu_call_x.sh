#!/bin/bash
# K0 of day d+1 on its own stream beside K1..K3 of day d: parity tests, then A/B (MTD_K0_PIPE=0 vs 1) at N=1
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/x_tests.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/x_tests.log
run() { MTD_K0_PIPE=$1 python bench.py --no-cpu-baseline $2 2>/dev/null | tee -a gpurun_out/x_bench_pipe$1.json | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('pipe=$1', round(j['ms_per_step'],4), (j.get('e2e') or {}).get('ms_per_step'), {k:round(v,3) for k,v in j['roofline']['stage_ms'].items()})"; }
rm -f gpurun_out/x_bench_pipe*.json
run 0 --no-e2e; run 1 --no-e2e; run 0 --no-e2e; run 1; 
