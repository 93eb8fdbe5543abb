#!/bin/bash
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:k1_ -s 42 -c 6 -o gpurun_out/h_k1_full -f \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/h_k1_full.log 2>&1; echo "ncu2=$?"
