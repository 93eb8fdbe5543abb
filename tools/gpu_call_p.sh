#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/p_gpu_tests.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/p_gpu_tests.log
M="dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_read_lookup_hit.sum"
timeout 500 ncu --metrics $M --clock-control none -k regex:k1_ -c 42 --csv --log-file gpurun_out/p_k1_new.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/p_ncu_new.log 2>&1; echo "new=$?"
for i in 1 2 3; do
( cd _wt_old && python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('old', j['ms_per_step'], j['roofline']['k1_ms'])" )
python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('new', j['ms_per_step'], j['roofline']['k1_ms'], j['roofline']['stage_ms'])"
done
