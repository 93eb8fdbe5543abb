#!/bin/bash
# host-side timing of the end-to-end loop at N > 1
mkdir -p gpurun_out
nproc > gpurun_out/k_probe.log
timeout 300 python tools/e2e_hostprobe.py --fake-world 8 >> gpurun_out/k_probe.log 2>&1; echo "fake8=$?"
timeout 300 python tools/e2e_hostprobe.py --fake-world 2 >> gpurun_out/k_probe.log 2>&1; echo "fake2=$?"
timeout 300 python tools/e2e_hostprobe.py --fake-world 8 --lag 0 >> gpurun_out/k_probe.log 2>&1; echo "fake8lag0=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/e2e_hostprobe.py >> gpurun_out/k_probe.log 2>&1; echo "n2=$?"
MTD_EXCHANGE=nccl timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/e2e_hostprobe.py >> gpurun_out/k_probe.log 2>&1; echo "n2nccl=$?"
grep -v "^\[" gpurun_out/k_probe.log | grep -E "rank|^[0-9]+$"
