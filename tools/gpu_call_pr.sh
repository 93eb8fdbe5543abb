#!/bin/bash
# main / state stream priority in the day pipeline: MTD_MAIN_PRIO=0 (default priority) vs high, at N=1
run() { env $1 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('$1', round(j['ms_per_step'],4), {k:round(v,3) for k,v in j['roofline']['stage_ms'].items()})"; }
run MTD_MAIN_PRIO=0; run MTD_MAIN_PRIO=1; run MTD_MAIN_PRIO=0; run MTD_MAIN_PRIO=1
