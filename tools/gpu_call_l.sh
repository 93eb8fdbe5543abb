#!/bin/bash
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/l_topo.log 2>&1
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 tools/pcie_multi.py > gpurun_out/l_pcie.log 2>&1; echo "rc=$?"
grep ranks gpurun_out/l_pcie.log; head -12 gpurun_out/l_topo.log
