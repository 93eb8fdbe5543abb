#!/bin/bash
# size-sorted small classes (MTD_K1_LPT bit 2): parity tests with the default (7), then A/B 3 vs 7 at N=1
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/t_tests.log 2>&1; echo "tests=$?"; tail -2 gpurun_out/t_tests.log
run() { MTD_K1_LPT=$1 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('N=1 lpt=$1', round(j['ms_per_step'],4), round(j['roofline']['k1_ms'],4))"; }
for rep in 1 2; do run 3; run 7; done
