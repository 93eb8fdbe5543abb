#!/bin/bash
# packed K1 attributes + per-slot activity column: parity, bench, and the DRAM traffic of the K1 launches of one day
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/n_gpu_tests.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/n_gpu_tests.log
python bench.py --no-cpu-baseline --no-e2e > gpurun_out/n_bench.json 2> gpurun_out/n_bench.err; echo "bench=$?"
python -c "
import json; j=json.load(open('gpurun_out/n_bench.json')); print(j['ms_per_step'], j['value'], j['roofline']['stage_ms'])"
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum --clock-control none \
    -k regex:k1_ -s 30 -c 6 --csv --log-file gpurun_out/n_k1_traffic.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/n_ncu.log 2>&1; echo "ncu=$?"
grep -c k1_ gpurun_out/n_k1_traffic.csv
