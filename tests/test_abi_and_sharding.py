"""CPU-side checks: the C-ABI library exports everything include/mtd_b200.h declares (no compute calls without a GPU),
and the host logic of the N>1 path -- the static cell -> rank owner map and the once-a-day exchange of the delta
slices (engine.exchange_deltas) -- over gloo with world_size 2."""
from __future__ import annotations

import ctypes
import os
import re
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mtdcovabm_b200 import capi, engine, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "mtd_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mtd_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = header_functions()
    assert len(names) >= 25
    lib = ctypes.CDLL(capi.SO)
    for n in names:
        assert hasattr(lib, n), "libmtd_b200.so does not export %s" % n
    assert sorted(capi.EXPORTS) == names, "capi.EXPORTS and include/mtd_b200.h disagree"
    lib.mtd_abi_version.restype = ctypes.c_int
    assert lib.mtd_abi_version() >= 1


def test_struct_layouts_match_header_constants():
    src = open(os.path.join(ROOT, "include", "mtd_b200.h")).read()
    assert int(re.search(r"#define MTD_DELTA_WORDS (\d+)", src).group(1)) == capi.DELTA_WORDS
    assert int(re.search(r"#define MTD_MAXSCHED (\d+)", src).group(1)) == capi.MAXSCHED
    assert ctypes.sizeof(capi.Counters) == 8 * (8 + 9 + 1 + 1)       # 8 scalars, class_cells[9], n_incidences, error
    assert int(re.search(r"#define MTD_MAX_CELL_INTERVALS (\d+)", src).group(1)) == capi.MAX_CELL_INTERVALS
    assert ctypes.sizeof(capi.Day) == 32


def test_missing_device_fails_loudly():
    """no CPU fallback: without a CUDA device the engine refuses to construct"""
    if torch.cuda.is_available():
        return
    try:
        engine.DayEngine(10, 10, 10, None)
    except RuntimeError as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("DayEngine constructed without a CUDA device")


def test_cell_owner_partitions_cells():
    cell = torch.arange(0, 200_000, dtype=torch.int32)
    for world in (1, 2, 4, 8):
        o = synth.cell_owner(cell, world)
        assert int(o.min()) >= 0 and int(o.max()) < world
        cnt = torch.bincount(o.to(torch.int64), minlength=world).double()
        assert float(cnt.max() / cnt.mean()) < 1.05          # cells spread evenly
        assert torch.equal(o, synth.cell_owner(cell, world))  # static


def test_balanced_owner_map_evens_out_the_expected_cell_cost():
    """greedy min-heap on the cells' expected cost (reference precedent util.py:675-707): with the
    bank's heavy-tailed cell sizes every rank's load is within 1 % of the mean, where the hash map is several % off"""
    bank = synth.load_bank()
    C1 = len(bank["cell_type"])
    w1 = np.zeros(C1)
    for k in range(int(bank["n_days"])):
        m = np.bincount(bank["d%d_iv_cell" % k].astype(np.int64), minlength=C1).astype(np.float64)
        w1 += m * m + 16.0 * m
    w = np.tile(w1, 64)
    for world in (2, 4, 8):
        o = synth.balanced_cell_owner(w, world)
        assert o.min() >= 0 and o.max() < world
        load = np.bincount(o, weights=w, minlength=world)
        assert load.max() / load.mean() < 1.01, (world, load / load.mean())
        assert np.array_equal(o, synth.balanced_cell_owner(w, world))       # deterministic: every rank computes the same map
        # small heavy-tailed case through the snake part as well
        o2 = synth.balanced_cell_owner(w[:5000], world, exact=100)
        l2 = np.bincount(o2, weights=w[:5000], minlength=world)
        assert l2.max() / l2.mean() < 1.10


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, cap, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # interval shard of this rank: every interval belongs to exactly one rank, whole cells stay together
    rng = np.random.default_rng(5)
    cell = torch.from_numpy(rng.integers(0, 5000, 60_000).astype(np.int32))
    mine = synth.cell_owner(cell, world) == rank
    n_mine = torch.tensor([int(mine.sum())])
    dist.all_reduce(n_mine)
    ok_cover = int(n_mine) == cell.numel()
    # the delta exchange: this rank's records (count in the header word) land in slice `rank` on every rank
    gather = torch.zeros((world, cap + 1, capi.DELTA_WORDS), dtype=torch.int32)
    n_rec = 3 + 2 * rank
    gather[rank, 0, 0] = n_rec
    gather[rank, 1:1 + n_rec] = (torch.arange(n_rec * capi.DELTA_WORDS, dtype=torch.int32)
                                 .reshape(n_rec, capi.DELTA_WORDS) + 1000 * (rank + 1))
    engine.exchange_deltas(gather, rank)
    ok = ok_cover
    for r in range(world):
        nr = 3 + 2 * r
        ok &= int(gather[r, 0, 0]) == nr
        ok &= bool((gather[r, 1:1 + nr].flatten() == torch.arange(nr * capi.DELTA_WORDS, dtype=torch.int32)
                    + 1000 * (r + 1)).all())
        ok &= bool((gather[r, 1 + nr:] == 0).all())
    # every rank must hold the identical gathered buffer (the apply step is deterministic given it)
    chk = gather.to(torch.int64).sum().reshape(1)
    lst = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(lst, chk)
    ok &= all(int(x) == int(chk) for x in lst)
    out[rank] = int(ok)
    dist.destroy_process_group()


def test_two_rank_delta_exchange_gloo():
    world = 2
    port = _free_port()
    out = mp.get_context("spawn").Manager().dict()
    mp.spawn(_worker, args=(world, port, 64, out), nprocs=world, join=True)
    assert dict(out) == {0: 1, 1: 1}


def test_package_import_sets_hardware_queue_count_for_multi_rank_runs_only():
    """mtdcovabm_b200/__init__.py: CUDA_DEVICE_MAX_CONNECTIONS=32 under WORLD_SIZE > 1 unless the user set a value"""
    import subprocess
    import sys
    code = "import os, mtdcovabm_b200; print(os.environ.get('CUDA_DEVICE_MAX_CONNECTIONS', 'unset'))"

    def run(extra):
        env = {k: v for k, v in os.environ.items() if k not in ("WORLD_SIZE", "CUDA_DEVICE_MAX_CONNECTIONS")}
        env.update(extra)
        return subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True).stdout.strip()

    assert run({}) == "unset"
    assert run({"WORLD_SIZE": "8"}) == "32"
    assert run({"WORLD_SIZE": "8", "CUDA_DEVICE_MAX_CONNECTIONS": "4"}) == "4"
