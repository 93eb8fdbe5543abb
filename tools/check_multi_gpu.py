#!/usr/bin/env python
"""Multi-GPU == single-GPU, bit for bit: every rank runs the cell-sharded engine (world = N, peer
exchange or NCCL) and, on the same GPU, the single-GPU engine over the whole population, for `--days` days, and compares
the replicated per-agent state after the last day and the per-day exposure counts.  Launch under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tools/check_multi_gpu.py --reps 40 --days 9 [--host]
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mtdcovabm_b200 import engine, params as prm, synth  # noqa: E402

KEYS = ("seir", "inf_type", "inf_sev", "sched", "sched_cnt", "test_day", "test_result_day", "quarantine", "hosp")


def run(eng, pop, streams, days, host, overlap=False):
    nb = len(streams)
    exposed = []
    if host:
        hs = [{k: (v.cpu().pin_memory() if v is not None else None) for k, v in s.items() if k != "n_total"} for s in streams]
        cap = max(4 * len(s["iv_agent"]) for s in streams) + 1024
        eng.enable_host_pipeline(cap)
        res = [eng.step_host(1 + d, hs[d % nb], next_s=hs[(d + 1) % nb], testing=True) for d in range(days)][-2:]
        # results of a day stay valid until day d+2 is submitted: only the last two can still be read
        rows = sum(r.n_rows for r in res)
        exposed = [r.n_exposed for r in res]
    elif overlap:
        # the bench's configuration: pre-built device streams, K0 of day d+1 beside K1..K3 of day d, no host read
        # between the days (the host runs ahead, so the stages really overlap)
        torch.cuda.synchronize()
        eng.set_inputs_ready(True)
        for d in range(days):
            eng.step(1 + d, streams[d % nb], testing=True)
        exposed = [int(eng.h.counters().n_exposed)]
        rows = int(eng.h.counters().n_rows)
    else:
        for d in range(days):
            eng.step(1 + d, streams[d % nb], testing=True)
            exposed.append(int(eng.h.counters().n_exposed))
        rows = int(eng.h.counters().n_rows)
    eng.sync()
    return eng.h.get_state(), exposed, rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=40)
    ap.add_argument("--days", type=int, default=9)
    ap.add_argument("--host", action="store_true", help="drive both engines through the pipelined host boundary")
    ap.add_argument("--overlap", action="store_true", help="the N-rank engine runs with mtd_set_k0_overlap on and no per-day host reads")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    bank = synth.load_bank()
    pop = synth.SynthPopulation(bank, a.reps, dev, seed=5, infectious_frac=0.05)
    P = prm.build_params()
    nb = pop.n_days
    full = [pop.stream(k, 0, 1) for k in range(nb)]
    mine = [pop.stream(k, rank, world) for k in range(nb)]
    max_m = max(len(s["iv_agent"]) for s in full) + 16

    e1 = engine.DayEngine(pop.n_agents, pop.n_cells, max_m, P, device=local, rank=0, world=1, seed=5)
    e1.set_population(pop)
    st1, ex1, rows1 = run(e1, pop, full, a.days, a.host)
    e1.close()

    eN = engine.DayEngine(pop.n_agents, pop.n_cells, max_m, P, device=local, rank=rank, world=world, seed=5)
    eN.set_population(pop)
    stN, exN, rowsN = run(eN, pop, mine, a.days, a.host, a.overlap)
    if a.overlap:
        ex1 = ex1[-1:]
    kind = eN.exchange
    eN.close()

    bad = [k for k in KEYS if not np.array_equal(st1[k], stN[k])]
    t = torch.tensor([rowsN], dtype=torch.int64, device=dev)
    dist.all_reduce(t)
    ok = not bad and ex1 == exN and int(t.item()) == rows1 and sum(ex1) > 0
    flag = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("multi-gpu check: world=%d exchange=%s days=%d agents=%d exposures/day=%s -> %s" %
              (world, kind, a.days, pop.n_present, ex1, "EQUAL" if int(flag.item()) else "DIFFERENT"), flush=True)
    if not ok:
        print("rank %d: differing fields %s, exposures %s vs %s, rows %d vs %d" % (rank, bad, ex1, exN, int(t.item()), rows1),
              flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
