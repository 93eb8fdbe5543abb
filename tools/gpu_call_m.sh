#!/bin/bash
# K9 seed exposures: new tests + the whole GPU suite + day-loop config run + bench sanity
mkdir -p gpurun_out
python -m pytest tests/test_seed_exposures.py tests/test_dayloop.py -m gpu -x -q > gpurun_out/m_seed_tests.log 2>&1; echo "seed_tests=$?"; tail -5 gpurun_out/m_seed_tests.log
python -m pytest tests -m gpu -x -q > gpurun_out/m_gpu_tests.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/m_gpu_tests.log
python tools/run_config.py --agents 5000000 --days 365 > gpurun_out/m_run5m.log 2>&1; echo "run5m=$?"; tail -4 gpurun_out/m_run5m.log
python bench.py --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/m_bench.json 2> gpurun_out/m_bench.err; echo "bench=$?"
python -c "
import json; j=json.load(open('gpurun_out/m_bench.json')); print(j['ms_per_step'], j['value'], j['e2e']['ms_per_step'])"
