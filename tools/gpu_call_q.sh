#!/bin/bash
# quick check of the serial roofline fields of bench.py
mkdir -p gpurun_out
python bench.py --no-cpu-baseline --no-e2e --steps 14 > gpurun_out/q_bench.json 2>/dev/null; echo rc=$?
python -c "
import json; j=json.load(open('gpurun_out/q_bench.json')); r=j['roofline']; print(j['ms_per_step'], r['frac'], r['k1_ms_serial'], r['achieved_serial'], r['frac_serial'])"
