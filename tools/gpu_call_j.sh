#!/bin/bash
# final evidence of the round: smoke, full GPU tests, default bench (incl. CPU baseline), reference arm, launch list
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/j_smoke.log 2>&1; echo "smoke=$?"; tail -1 gpurun_out/j_smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/j_gpu_tests.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/j_gpu_tests.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/j_bench_ref.json 2> gpurun_out/j_bench_ref.err; echo "ref=$?"
python bench.py > gpurun_out/j_bench.json 2> gpurun_out/j_bench.err; echo "bench=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 340 --csv --log-file gpurun_out/j_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/j_ncu.log 2>&1; echo "ncu1=$?"
