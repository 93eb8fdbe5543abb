#!/bin/bash
# before/after of the packed K1 attributes: DRAM traffic + time of the K1 launches of the 7 calibration days (42 launches)
mkdir -p gpurun_out
M="dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_read_lookup_hit.sum"
( cd _wt_old && timeout 500 ncu --metrics $M --clock-control none -k regex:k1_ -c 42 --csv --log-file ../gpurun_out/o_k1_old.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > ../gpurun_out/o_ncu_old.log 2>&1; echo "old=$?" )
timeout 500 ncu --metrics $M --clock-control none -k regex:k1_ -c 42 --csv --log-file gpurun_out/o_k1_new.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/o_ncu_new.log 2>&1; echo "new=$?"
for i in 1 2 3; do
( cd _wt_old && python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('old', j['ms_per_step'], j['roofline']['k1_ms'])" )
python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('new', j['ms_per_step'], j['roofline']['k1_ms'])"
done
