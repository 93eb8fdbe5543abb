#!/bin/bash
# does stream aliasing onto the default 8 hardware queues serialise the day pipeline?  CUDA_DEVICE_MAX_CONNECTIONS sweep
run() { env $1 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('$1', round(j['ms_per_step'],4), {k:round(v,3) for k,v in j['roofline']['stage_ms'].items()})"; }
run CUDA_DEVICE_MAX_CONNECTIONS=8; run CUDA_DEVICE_MAX_CONNECTIONS=16; run CUDA_DEVICE_MAX_CONNECTIONS=32
run "CUDA_DEVICE_MAX_CONNECTIONS=32 MTD_K1_PIPE=0"
