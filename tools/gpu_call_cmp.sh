#!/bin/bash
# config 3 (520k agents x 180 days, every stage on) at HEAD and at the commit in _wt_old
for i in 1 2; do
python tools/run_config.py --agents 520000 --days 180 --params epidemiology.json --every 1000 2>&1 | tail -1 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('head', j['wall_s'], j['ms_per_day'])"
( cd _wt_old && python tools/run_config.py --agents 520000 --days 180 --params epidemiology.json --every 1000 2>&1 | tail -1 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('old ', j['wall_s'], j['ms_per_day'])" )
done
