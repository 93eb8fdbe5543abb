#!/bin/bash
mkdir -p gpurun_out
for n in 8 4; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n \
    > gpurun_out/e_bench_n$n.json 2> gpurun_out/e_bench_n$n.err; echo "bench$n=$?"
done
