#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/e_gpu_tests.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/e_gpu_tests.log
python bench.py --no-cpu-baseline > gpurun_out/e_bench_n1.json 2> gpurun_out/e_bench_n1.err; echo "bench1=$?"
for n in 8 4 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n \
    > gpurun_out/e_bench_n$n.json 2> gpurun_out/e_bench_n$n.err; echo "bench$n=$?"
done
