#!/usr/bin/env python
"""Run one of the tiled throughput configurations end to end with every stage of the state step switched on
(mtdcovabm_b200/dayloop.py):   python tools/run_config.py --agents 520000 --days 180 [--params epidemiology_strict_days.json]
On N GPUs (cells sharded over the ranks, BASELINE.json configs[3] "5M agents cell-sharded across 8xB200, 365 days"):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29555 \
        tools/run_config.py --agents 5000000 --days 365
Prints a per-day line every `--every` days and a JSON summary (rank 0); exits non-zero on any device error / overflow."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mtdcovabm_b200 import dayloop, params as prm, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--agents", type=int, default=520_000)
    ap.add_argument("--days", type=int, default=180)
    ap.add_argument("--params", default="epidemiology_strict_days.json")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--every", type=int, default=15)
    a = ap.parse_args()
    np.random.seed(a.seed)
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=dev)
    bank = synth.load_bank()
    reps = max(1, round(a.agents / int(bank["present"].sum())))
    pop = synth.SynthPopulation(bank, reps, dev, seed=a.seed, infectious_frac=0.002)
    run = dayloop.EpidemicRun(pop, prm.load_json(a.params), device=local, seed=a.seed, rank=rank, world=world)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for day in range(1, a.days + 1):
        r = run.day(day)
        if rank == 0 and (day % a.every == 0 or day == 1):
            print("day %3d  S/E/I/R/D %s  exposed %6d  rows %9d  traced %5d (+%d primary, %d secondary)  vacc %6d  mhd %.2f"
                  % (day, r["occupancy"][1:], r["exposed"], r["rows"], r["traced"], r["primary"], r["secondary"],
                     r["vaccinations"], r["mhd"]), flush=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    log = run.log
    summ = {"n_gpus": world, "exchange": run.exchange, "agents": run.n_present, "days": a.days, "params": a.params, "wall_s": dt, "ms_per_day": 1e3 * dt / a.days,
            "agent_days_per_s": run.n_present * a.days / dt, "pairs_per_s": sum(r["pairs"] for r in log) / dt,
            "total_exposed": sum(r["exposed"] for r in log), "total_traced": sum(r["traced"] for r in log),
            "total_primary": sum(r["primary"] for r in log), "total_vaccinations": sum(r["vaccinations"] for r in log),
            "peak_infectious": max(r["occupancy"][3] for r in log), "deceased": log[-1]["occupancy"][5],
            "sched_overflow": sum(r["sched_overflow"] for r in log)}
    if rank == 0:
        print(json.dumps(summ), flush=True)
    run.close()
    if world > 1:
        torch.distributed.destroy_process_group()
    sys.exit(1 if summ["sched_overflow"] else 0)


if __name__ == "__main__":
    main()
