#!/usr/bin/env python
"""The WHOLE state step on N GPUs == on one GPU, bit for bit: every rank runs dayloop.EpidemicRun cell-sharded
(world = N: K3 exposure records over the peer exchange, K5 primary-contact flags over one all-reduce, K4 / K6 / K7 on the
replicated state) and, on the same GPU, the single-GPU run over the whole population, for `--days` days under the
strict-interventions parameter file, and compares every per-agent record, the vaccination records and the per-day log
(exposed, traced, primary / secondary contacts, vaccinations, occupancy).  Launch under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 \
        tools/check_multi_gpu_dayloop.py --reps 20 --days 40
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

from mtdcovabm_b200 import dayloop, params as prm, synth  # noqa: E402

LOG_KEYS = ("occupancy", "pairs", "rows", "exposed", "traced", "primary", "secondary", "vaccinations", "sched_overflow",
            "mhd", "testing", "quarantine", "tracing")


def run(pop, ep, local, seed, days, rank, world):
    r = dayloop.EpidemicRun(pop, ep, device=local, seed=seed, rank=rank, world=world)
    log = [r.day(d) for d in range(1, days + 1)]
    st = r.h.get_state()
    last, rec = r.h.get_vaccination()
    kind = r.exchange
    r.close()
    return log, st, last, rec, kind


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--days", type=int, default=40)
    ap.add_argument("--seed", type=int, default=3)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    bank = synth.load_bank()
    pop = synth.SynthPopulation(bank, a.reps, dev, seed=a.seed, infectious_frac=0.01)
    ep = dict(prm.load_json("epidemiology_strict_days.json"))
    ep["vaccination_day_thresholds"] = [[0, 0.0], [10, 0.4]]
    ep["vaccination_infectiousrate_thresholds"] = []
    log1, st1, last1, rec1, _ = run(pop, ep, local, a.seed, a.days, 0, 1)
    logN, stN, lastN, recN, kind = run(pop, ep, local, a.seed, a.days, rank, world)
    bad = [k for k in st1 if not np.array_equal(st1[k], stN[k])]
    if not np.array_equal(last1, lastN):
        bad.append("last_dose_day")
    if not np.array_equal(rec1, recN):
        bad.append("vacc_day")
    for d, (r1, rN) in enumerate(zip(log1, logN)):
        for k in LOG_KEYS:
            if r1[k] != rN[k]:
                bad.append("day %d %s: %s vs %s" % (d + 1, k, r1[k], rN[k]))
    work = (sum(r["exposed"] for r in log1) > 50 and sum(r["traced"] for r in log1) > 0 and sum(r["primary"] for r in log1) > 0
            and sum(r["vaccinations"] for r in log1) > 0)
    ok = not bad and work
    flag = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("multi-gpu day loop: world=%d exchange=%s days=%d agents=%d exposed=%d traced=%d primary=%d secondary=%d "
              "vaccinations=%d -> %s" % (world, kind, a.days, pop.n_present, sum(r["exposed"] for r in log1),
                                         sum(r["traced"] for r in log1), sum(r["primary"] for r in log1),
                                         sum(r["secondary"] for r in log1), sum(r["vaccinations"] for r in log1),
                                         "EQUAL" if int(flag.item()) else "DIFFERENT"), flush=True)
    if not ok:
        print("rank %d: %s%s" % (rank, bad[:12], "" if work else " (no work done)"), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
