#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/c_gpu_tests.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/c_gpu_tests.log
python tools/e2e_probe.py > gpurun_out/c_e2e_probe.txt 2>&1; echo "probe=$?"
python bench.py --no-cpu-baseline --no-e2e > gpurun_out/c_bench_n1.json 2> gpurun_out/c_bench_n1.err; echo "bench1=$?"
