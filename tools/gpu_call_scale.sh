#!/bin/bash
# bench at N GPUs of one box (N = $1); at N = 2 also the multi-GPU parity tests
N=${1:-2}
mkdir -p gpurun_out
if [ "$N" = 2 ]; then python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/s_tests_n2.log 2>&1; echo "tests=$?"; tail -2 gpurun_out/s_tests_n2.log; fi
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N \
    > gpurun_out/s_bench_n$N.json 2> gpurun_out/s_bench_n$N.err; echo "bench=$?"
python -c "
import json; j=json.load(open('gpurun_out/s_bench_n$N.json')); print('N=$N', j['ms_per_step'], j['value'], (j.get('e2e') or {}).get('ms_per_step'), j['roofline']['stage_ms'])"
