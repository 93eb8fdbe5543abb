#!/bin/bash
# HEAD: 7-day K1 traffic/instruction capture + one full (source-level) capture of a day's K1 launches
mkdir -p gpurun_out
M="dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_read_lookup_hit.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active"
timeout 600 ncu --metrics $M --clock-control none -k regex:k1_ -c 70 --csv --log-file gpurun_out/w_k1_traffic.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/w_ncu.log 2>&1; echo "ncu=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k1_ -s 10 -c 10 -o gpurun_out/w_k1_full -f \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/w_k1_full.log 2>&1; echo "ncu2=$?"
ls -la gpurun_out/
