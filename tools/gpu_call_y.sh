#!/bin/bash
# K0 overlap: stream priority (0 low, 1 high, 2 default) and staged-input overlap in the e2e path
mkdir -p gpurun_out
run() { env $1 python bench.py --no-cpu-baseline $2 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('$1', round(j['ms_per_step'],4), (j.get('e2e') or {}).get('ms_per_step'), {k:round(v,3) for k,v in j['roofline']['stage_ms'].items()})"; }
run MTD_K0_PRIO=0 --no-e2e; run MTD_K0_PRIO=1 --no-e2e; run MTD_K0_PRIO=2 --no-e2e
run MTD_K0_PIPE_STAGED=0; run MTD_K0_PIPE_STAGED=1; run "MTD_K0_PIPE_STAGED=1 MTD_K0_PRIO=1"
