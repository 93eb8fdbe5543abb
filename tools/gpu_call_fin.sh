#!/bin/bash
# final check at the round's last commit: smoke, GPU tests, default bench at N=1
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/fin_smoke.log 2>&1; echo "smoke=$?"; tail -1 gpurun_out/fin_smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/fin_tests.log 2>&1; echo "tests=$?"; tail -2 gpurun_out/fin_tests.log
python bench.py > gpurun_out/fin_bench.json 2> gpurun_out/fin_bench.err; echo "bench=$?"; tail -3 gpurun_out/fin_bench.err
python -c "
import json; j=json.load(open('gpurun_out/fin_bench.json')); r=j['roofline']; print(j['ms_per_step'], j['e2e']['ms_per_step'], r['stage_ms'], r['stage_ms_serial'], j['gpu_launches'], j['cpu_baseline'])"
