#!/bin/bash
# evidence of HEAD: smoke, full GPU tests, reference arm, default bench (CPU baseline + e2e), launch list, K1 traffic/instructions
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v_smoke.log 2>&1; echo "smoke=$?"; tail -1 gpurun_out/v_smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/v_gpu_tests.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/v_gpu_tests.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/v_bench_ref.json 2> gpurun_out/v_bench_ref.err; echo "ref=$?"
python bench.py > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; echo "bench=$?"
python -c "
import json; j=json.load(open('gpurun_out/v_bench.json')); print(j['ms_per_step'], j['value'], j['e2e']['ms_per_step'], j['roofline']['stage_ms'], j['cpu_baseline'])"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 340 --csv --log-file gpurun_out/v_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/v_ncu.log 2>&1; echo "ncu1=$?"
M="dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active"
timeout 600 ncu --metrics $M --clock-control none -k regex:k1_ -c 42 --csv --log-file gpurun_out/v_k1_traffic.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/v_ncu2.log 2>&1; echo "ncu2=$?"
