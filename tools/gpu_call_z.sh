#!/bin/bash
# K1 small: three-deep prefetch pipeline + branch-free attribute gather: parity tests, then stage times (serial K0 for a clean k1_ms; then overlapped)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/z_tests.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/z_tests.log
run() { env $1 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('$1', round(j['ms_per_step'],4), {k:round(v,3) for k,v in j['roofline']['stage_ms'].items()})"; }
run MTD_K0_PIPE=0; run MTD_K0_PIPE=0
