"""Cell-sharded multi-GPU day step == single-GPU day step, bit for bit: tools/check_multi_gpu.py
under torchrun on 2 GPUs, through the device-resident API and through the pipelined host boundary, over the NVLink
peer exchange and over the NCCL all-gather.  Skipped on a single-GPU box (the 2-rank host logic is covered on CPU over
gloo in test_abi_and_sharding.py)."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _run(extra, env_extra=None, port=29541):
    env = dict(os.environ)
    env.update(env_extra or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tools", "check_multi_gpu.py"), "--reps", "40",
           "--days", "9"] + extra
    p = subprocess.run(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-4000:]
    assert "EQUAL" in p.stdout, p.stdout[-4000:]
    return p.stdout


needs2 = pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")


@needs2
def test_two_gpus_equal_one_gpu_peer_exchange():
    out = _run([], port=29541)
    assert "exchange=p2p" in out or "exchange=nccl" in out


@needs2
def test_two_gpus_equal_one_gpu_nccl_allgather():
    out = _run([], {"MTD_EXCHANGE": "nccl"}, port=29542)
    assert "exchange=nccl" in out


@needs2
def test_two_gpus_equal_one_gpu_host_boundary():
    _run(["--host"], port=29543)


@needs2
def test_two_gpus_with_k0_overlap_equal_one_gpu_serial():
    """the bench's N > 1 configuration: K0 of day d+1 beside K1..K3 of day d on every rank, peer exchange, host running ahead"""
    _run(["--overlap"], port=29544)


@needs2
def test_two_gpu_whole_state_step_equals_one_gpu():
    """40 days of dayloop.EpidemicRun (K4, K6, K0 with the intervention mask, K1-K3, K5 over the sharded row ring, K7) on
    2 GPUs == on 1 GPU: every per-agent record, the vaccination records and the per-day log, bit for bit"""
    env = dict(os.environ)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29545", os.path.join(ROOT, "tools", "check_multi_gpu_dayloop.py"), "--reps", "20",
           "--days", "40"]
    p = subprocess.run(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1200)
    assert p.returncode == 0, p.stdout[-4000:]
    assert "EQUAL" in p.stdout, p.stdout[-4000:]
