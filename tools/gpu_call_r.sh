#!/bin/bash
# longest-first order of the 129-320 class: A/B at N=4 and N=1 (device-resident loop only)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_edge_cases.py tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/r_tests.log 2>&1; echo "tests=$?"; tail -2 gpurun_out/r_tests.log
run() { # $1 = N, $2 = lpt
  if [ "$1" = 1 ]; then MTD_K1_LPT=$2 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null
  else MTD_K1_LPT=$2 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 295$1$2 bench.py --gpus $1 --no-e2e 2>/dev/null; fi | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('N=$1 lpt=$2', round(j['ms_per_step'],4), round(j['roofline']['k1_ms'],4))"
}
for rep in 1 2; do run 4 0; run 4 1; done
for rep in 1 2; do run 1 0; run 1 1; done
