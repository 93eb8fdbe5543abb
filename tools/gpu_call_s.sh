#!/bin/bash
mkdir -p gpurun_out
run() { MTD_K1_LPT=$1 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('N=1 lpt=$1', round(j['ms_per_step'],4), round(j['roofline']['k1_ms'],4))"; }
for rep in 1 2; do run 0; run 1; run 3; done
MTD_K1_LPT=3 python -m pytest tests/test_gpu_parity.py tests/test_edge_cases.py -m gpu -x -q 2>&1 | tail -1
