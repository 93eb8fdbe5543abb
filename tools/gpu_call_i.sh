#!/bin/bash
mkdir -p gpurun_out
( echo "### config 2 shape: 10k agents, epidemiology_normal_days.json, 90 days"; python tools/run_config.py --agents 10000 --days 90 --params epidemiology_normal_days.json --every 30
  echo "### config 5 shape: 10k agents, epidemiology_strict_days.json (testing + quarantine + contact tracing), 90 days"; python tools/run_config.py --agents 10000 --days 90 --params epidemiology_strict_days.json --every 30
  echo "### config 3: 520k agents, epidemiology.json, 180 days"; python tools/run_config.py --agents 520000 --days 180 --params epidemiology.json --every 30
  echo "### config 4 (one GPU): 5M agents, epidemiology.json, 365 days"; python tools/run_config.py --agents 5000000 --days 365 --params epidemiology.json --every 60
  echo "### config 4 with strict interventions: 5M agents, 365 days"; python tools/run_config.py --agents 5000000 --days 365 --params epidemiology_strict_days.json --every 60
) > gpurun_out/i_config_runs.txt 2>&1; echo "runs=$?"
