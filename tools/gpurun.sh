#!/bin/bash
# One entry point for the GPU-box calls of this repo (run ON the box through gpurun):
#   /usr/local/graft/bin/gpurun --timeout 900 -- 'bash tools/gpurun.sh <recipe> [tag]'
# Every recipe writes its logs under gpurun_out/<tag>_* (tag defaults to the recipe name).
#   quick     the K1-facing parity tests (edge cases, golden days, drop-in) + a short bench
#   tests     smoke + the whole `-m gpu` suite
#   bench     default bench.py at N=1 (+ reference arm)
#   scale N   bench.py on N GPUs through torchrun (call gpurun with --gpus N) + the multi-GPU tests
#   launches  ncu launch list (gpu__time_duration) of a short bench
#   ncu_k1    ncu --set full of the K1 kernels of one day (+ metrics csv over the seven day streams)
#   perday    per-day-stream stage times (tools/k1_per_day.py)
#   dictapi   cost of the reference-shaped dict API on the golden days (tools/time_dict_api.py)
#   configs   the BASELINE.json configuration shapes through the whole day loop (tools/run_config.py)
# A/B of environment switches on the device-resident bench: tools/ab.sh (one GPU; BENCH_ARGS="--shard-of 8" runs one rank's
# shard of an 8-rank job alone) and tools/ab_n.sh (N GPUs through torchrun); their lines are collected in
# profiles/r02_k1_schedule_ab.txt.
set -u
R=${1:-quick}; T=${2:-$R}; O=gpurun_out; mkdir -p $O
case $R in
quick)
  # both K1 implementations (MTD_K1_IMPL=block | legacy): parity tests, then a short bench each
  for impl in ${IMPLS:-split block legacy}; do
    MTD_K1_IMPL=$impl timeout 1500 python -m pytest tests/test_edge_cases.py tests/test_gpu_parity.py tests/test_dropin_api.py tests/test_k6_kat.py -m gpu -q > $O/${T}_tests_$impl.log 2>&1; echo "tests[$impl]=$?"; tail -n 12 $O/${T}_tests_$impl.log
    MTD_K1_IMPL=$impl timeout 600 python bench.py --steps 14 --warmup 7 --no-cpu-baseline --no-e2e > $O/${T}_bench_$impl.json 2> $O/${T}_bench_$impl.err; echo "bench[$impl]=$?"; tail -n 3 $O/${T}_bench_$impl.err
    python -c "
import json; j=json.load(open('$O/${T}_bench_$impl.json')); r=j['roofline']; print('$impl', j['ms_per_step'], r['stage_ms'], r['stage_ms_serial'])"
  done
  ;;
tests)
  python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke=$?"; tail -n 2 $O/${T}_smoke.log
  timeout 3000 python -m pytest tests -m gpu -x -q > $O/${T}_tests.log 2>&1; echo "tests=$?"; tail -n 15 $O/${T}_tests.log
  ;;
bench)
  timeout 900 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench=$?"; tail -n 3 $O/${T}_bench.err; cat $O/${T}_bench.json
  timeout 600 python bench.py --impl reference > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; echo "ref=$?"; cat $O/${T}_bench_ref.json
  ;;
scale)
  N=${3:-2}
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N \
      > $O/${T}_bench_n$N.json 2> $O/${T}_bench_n$N.err; echo "bench=$?"; tail -n 3 $O/${T}_bench_n$N.err; cat $O/${T}_bench_n$N.json
  timeout 1200 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > $O/${T}_tests_n$N.log 2>&1; echo "tests=$?"; tail -n 5 $O/${T}_tests_n$N.log
  ;;
launches)
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $O/${T}_launches.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > $O/${T}_ncu.log 2>&1; echo "ncu=$?"
  python tools/launch_summary.py $O/${T}_launches.csv | tee $O/${T}_launches_summary.txt
  ;;
ncu_k1)
  M="dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_read_lookup_hit.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active"
  timeout 900 ncu --metrics $M --clock-control none -k regex:k1_ -c 120 --csv --log-file $O/${T}_traffic.csv \
      python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > $O/${T}_ncu.log 2>&1; echo "ncu=$?"
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k1_ -s 24 -c 12 -o $O/${T}_full -f \
      python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > $O/${T}_full.log 2>&1; echo "ncu2=$?"
  ls -la $O/ | tail -n 5
  ;;
perday)
  python tools/k1_per_day.py 2>&1 | tail -n 16 | tee $O/${T}_perday.txt
  ;;
dictapi)
  timeout 900 python tools/time_dict_api.py > $O/${T}_dict_api_cost.json 2> $O/${T}_dict_api.err; echo "dictapi=$?"; tail -n 25 $O/${T}_dict_api.err
  ;;
configs)
  timeout 1500 python tools/run_config.py 2>&1 | tee $O/${T}_runs.txt | tail -n 30
  ;;
*) echo "unknown recipe $R"; exit 2;;
esac
