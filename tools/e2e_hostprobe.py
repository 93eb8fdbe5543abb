#!/usr/bin/env python
"""Where does the HOST spend the end-to-end day at N > 1?  Runs bench.py's step_host loop with a wall-clock timer around
every library call and prints the per-call mean per rank.

    torchrun --nproc-per-node 2 tools/e2e_hostprobe.py            # the N-rank loop with the peer exchange
    python tools/e2e_hostprobe.py --fake-world 8                     # ONE rank stepping a 1/8 shard, no exchange
"""
from __future__ import annotations

import argparse
import collections
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mtdcovabm_b200 import engine, params as prm, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--agents", type=int, default=5_000_000)
    ap.add_argument("--fake-world", type=int, default=0)
    ap.add_argument("--lag", type=int, default=1)
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    bank = synth.load_bank()
    reps = max(1, round(a.agents / int(bank["present"].sum())))
    pop = synth.SynthPopulation(bank, reps, dev)
    nb = pop.n_days
    sw = a.fake_world if a.fake_world else world
    streams = [pop.stream(k, rank, sw) for k in range(nb)]
    host = [{k: (v.cpu().pin_memory() if v is not None else None) for k, v in s.items() if k != "n_total"} for s in streams]
    max_m = max(len(s["iv_agent"]) for s in streams)
    eng = engine.DayEngine(pop.n_agents, pop.n_cells, max_m + 16, prm.build_params(), device=local, rank=rank, world=world)
    eng.set_population(pop)
    eng.enable_host_pipeline(4 * max_m + 1024)
    h = eng.h

    acc = collections.defaultdict(float)

    def timed(name):
        f = getattr(h, name)

        def g(*args, **kw):
            t = time.perf_counter()
            r = f(*args, **kw)
            acc[name] += time.perf_counter() - t
            return r
        setattr(h, name, g)

    for nm in ("advance_seir", "simulate_contact_network", "prefetch_intervals", "simulate_direct_contacts",
               "export_day_async", "export_wait", "make_day"):
        timed(nm)

    def run(first, count):
        pend = collections.deque()
        for step in range(first, first + count):
            res = eng.step_host(1 + step, host[step % nb], next_s=host[(step + 1) % nb])
            pend.append(res)
            if len(pend) > a.lag:
                pend.popleft().wait()
        while pend:
            pend.popleft().wait()

    run(0, 5)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    acc.clear()
    t0 = time.perf_counter()
    run(5, a.steps)
    torch.cuda.synchronize()
    tot = (time.perf_counter() - t0) * 1e3 / a.steps
    msg = "rank %d/%d shard 1/%d lag %d: %.3f ms/day | " % (rank, world, sw, a.lag, tot)
    msg += " ".join("%s=%.3f" % (k, v * 1e3 / a.steps) for k, v in sorted(acc.items()))
    print(msg, flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
