#!/bin/bash
# one GPU-box call: PCIe probe, default bench, launch list, full ncu capture of the K1 launches of one day
set -x
mkdir -p gpurun_out
nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/pcie_probe.cu -o /tmp/pcie_probe && /tmp/pcie_probe > gpurun_out/a_pcie.txt 2>&1
python bench.py > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; echo "bench=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/a_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/a_ncu.log 2>&1; echo "ncu1=$?"
ncu --set full --clock-control none --import-source on -k regex:k1_ -s 42 -c 6 -o gpurun_out/a_k1_full -f \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/a_k1_full.log 2>&1; echo "ncu2=$?"
