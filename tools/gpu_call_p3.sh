#!/bin/bash
# three-stage day pipeline (K0 | K1 | K2+K3 on three streams): parity tests, then A/B of MTD_K1_PIPE at N (default 1)
N=${1:-1}
mkdir -p gpurun_out
if [ "$N" = 1 ]; then python -m pytest tests -m gpu -x -q > gpurun_out/p3_tests.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/p3_tests.log
else python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/p3_tests_n$N.log 2>&1; echo "tests=$?"; tail -3 gpurun_out/p3_tests_n$N.log; fi
run() {
  if [ "$N" = 1 ]; then env $1 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null
  else env $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$N bench.py --gpus $N --no-e2e 2>/dev/null; fi | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('N=$N $1', round(j['ms_per_step'],4), {k:round(v,3) for k,v in j['roofline']['stage_ms'].items()})"
}
run MTD_K1_PIPE=0; run MTD_K1_PIPE=1; run "MTD_K0_PIPE=0"
