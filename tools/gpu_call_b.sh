#!/bin/bash
# 2-GPU box: GPU tests, multi-GPU equality, bench at N=1 and N=2 (peer exchange and NCCL)
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/b_gpu_tests.log 2>&1; echo "tests=$?"
tail -5 gpurun_out/b_gpu_tests.log
python bench.py --no-cpu-baseline > gpurun_out/b_bench_n1.json 2> gpurun_out/b_bench_n1.err; echo "bench1=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 \
    > gpurun_out/b_bench_n2.json 2> gpurun_out/b_bench_n2.err; echo "bench2=$?"
MTD_EXCHANGE=nccl python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 \
    > gpurun_out/b_bench_n2_nccl.json 2> gpurun_out/b_bench_n2_nccl.err; echo "bench2n=$?"
