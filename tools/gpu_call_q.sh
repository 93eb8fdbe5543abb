#!/bin/bash
# 2 GPUs: full GPU suite incl. the multi-GPU tests, N=1 and N=2 bench lines
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/q_gpu_tests.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/q_gpu_tests.log
python bench.py --no-cpu-baseline > gpurun_out/q_bench_n1.json 2> gpurun_out/q_bench_n1.err; echo "bench1=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 > gpurun_out/q_bench_n2.json 2> gpurun_out/q_bench_n2.err; echo "bench2=$?"
python -c "
import json
for n in (1,2):
    j=json.load(open('gpurun_out/q_bench_n%d.json'%n)); print(n, j['ms_per_step'], j['value'], j['e2e']['ms_per_step'], j['roofline'].get('sm_side'), j['roofline']['traffic'])"
