#!/usr/bin/env python
"""Do the ranks of one box share a host-side copy limit?  Every rank runs H2D and D2H copies of pinned buffers on two
streams at once; first one rank at a time, then groups of 2, 4, ... ranks together.
    torchrun --nproc-per-node 4 tools/pcie_multi.py"""
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
dist.init_process_group("gloo")
MB = 128
h_in = torch.empty(MB << 20, dtype=torch.uint8).pin_memory(); h_out = torch.empty(MB << 20, dtype=torch.uint8).pin_memory()
d_in = torch.empty(MB << 20, dtype=torch.uint8, device="cuda"); d_out = torch.empty(MB << 20, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def burst(reps, up, down):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if up:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if down:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


burst(3, True, True)
g = 1
while g <= world:
    for up, down in ((True, False), (False, True), (True, True)):
        dist.barrier()
        active = rank < g
        t = burst(20, up, down) if active else 0.0
        dist.barrier()
        res = [None] * world
        dist.all_gather_object(res, t)
        if rank == 0:
            per = [(MB / 1024 * (int(up) + int(down))) / x for x in res[:g]]
            print("ranks active %d  up=%d down=%d  per-rank GB/s %s  aggregate %.1f" % (
                g, up, down, " ".join("%.1f" % p for p in per), sum(per)), flush=True)
    g *= 2
dist.destroy_process_group()
