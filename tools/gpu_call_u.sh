#!/bin/bash
# K2 with the large cells taken from their class lists + lone-interval early-out in K1 small: parity, then stage times
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/u_tests.log 2>&1; echo "tests=$?"; tail -2 gpurun_out/u_tests.log
N=${1:-1}
run() {
  if [ "$N" = 1 ]; then python bench.py --no-cpu-baseline --no-e2e 2>/dev/null
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --no-e2e 2>/dev/null; fi | tee -a gpurun_out/u_bench_n$N.json | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('N=$N', round(j['ms_per_step'],4), {k:round(v,3) for k,v in j['roofline']['stage_ms'].items()})"
}
rm -f gpurun_out/u_bench_n$N.json
run; run
