#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/f_gpu_tests.log 2>&1; echo "tests=$?"; tail -15 gpurun_out/f_gpu_tests.log
python bench.py --no-cpu-baseline --no-e2e > gpurun_out/f_bench_n1.json 2> gpurun_out/f_bench_n1.err; echo "bench1=$?"
