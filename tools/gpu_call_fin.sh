#!/bin/bash
# final check of bench.py (serial stage pass added): default run at N=1
mkdir -p gpurun_out
python bench.py > gpurun_out/fin_bench.json 2> gpurun_out/fin_bench.err; echo "bench=$?"; tail -3 gpurun_out/fin_bench.err
python -c "
import json; j=json.load(open('gpurun_out/fin_bench.json')); r=j['roofline']; print(j['ms_per_step'], j['e2e']['ms_per_step'], r['stage_ms'], r['stage_ms_serial'], j['gpu_launches'], j['cpu_baseline'])"
