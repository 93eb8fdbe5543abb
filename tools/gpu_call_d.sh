#!/bin/bash
mkdir -p gpurun_out
python tools/e2e_probe.py --steps 20 > gpurun_out/d_e2e_probe3.txt 2>&1
python -m pytest tests -m gpu -x -q -k "host or pipel or export" > gpurun_out/d_tests.log 2>&1; tail -3 gpurun_out/d_tests.log
