#!/bin/bash
# the BASELINE.json configuration shapes through the full day loop (every stage on) at the round's last commit
mkdir -p gpurun_out
O=gpurun_out/cfg_runs.txt; : > $O
r() { echo "### $1" >> $O; shift; python tools/run_config.py "$@" >> $O 2>&1; echo "rc=$?"; }
r "config 2 shape: 10k agents, epidemiology_normal_days.json, 90 days" --agents 10000 --days 90 --params epidemiology_normal_days.json --every 30
r "config 5 shape: 10k agents, epidemiology_strict_days.json, 90 days" --agents 10000 --days 90 --params epidemiology_strict_days.json --every 30
r "config 3: 520k agents, epidemiology.json, 180 days" --agents 520000 --days 180 --params epidemiology.json --every 60
r "config 4 (one GPU): 5M agents, epidemiology.json, 365 days" --agents 5000000 --days 365 --params epidemiology.json --every 73
r "config 4 with strict interventions: 5M agents, 365 days" --agents 5000000 --days 365 --params epidemiology_strict_days.json --every 73
tail -3 $O | cut -c1-600
