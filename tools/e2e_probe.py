#!/usr/bin/env python
"""Where does the host boundary's time go?  Times the day loop (5M agents) with the upload and the export switched on
and off independently:  python tools/e2e_probe.py [--steps 30] [--agents 5000000]"""
from __future__ import annotations

import argparse
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mtdcovabm_b200 import capi, engine, params as prm, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--agents", type=int, default=5_000_000)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    bank = synth.load_bank()
    reps = max(1, round(a.agents / int(bank["present"].sum())))
    pop = synth.SynthPopulation(bank, reps, dev)
    nb = pop.n_days
    streams = [pop.stream(k) for k in range(nb)]
    host = [{k: (v.cpu().pin_memory() if v is not None else None) for k, v in s.items() if k != "n_total"} for s in streams]
    max_m = max(len(s["iv_agent"]) for s in streams)
    P = prm.build_params()
    eng = engine.DayEngine(pop.n_agents, pop.n_cells, max_m + 16, P, device=0)
    eng.set_population(pop)
    eng.enable_host_pipeline(4 * max_m)
    h = eng.h

    dev_rows = torch.empty((4 * max_m, 3), dtype=torch.int32, device=dev)
    dev_kv = torch.empty((eng._cap_exposed, 2), dtype=torch.int64, device=dev)
    dev_res = torch.zeros(capi.RESULT_WORDS, dtype=torch.int64, device=dev)

    def loop(upload, export, wait_mode):
        eng.set_population(pop)
        h.set_export(export)
        eng._staged = None
        prev = None
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for step in range(a.steps):
            d = h.make_day(capi.MODE_PHILOX, 1 + step, 0.0, False, False, seed=0)
            s, hs, nx = streams[step % nb], host[step % nb], host[(step + 1) % nb]
            with torch.cuda.stream(eng.stream):
                h.advance_seir(1 + step, None)
                if not upload:
                    h.simulate_contact_network(d, s["iv_agent"], s["iv_cell"], s["iv_start"], s["iv_end"], mem=capi.MEM_DEVICE)
                elif eng._staged is hs:
                    h.simulate_contact_network(d, None, None, None, None, n=len(hs["iv_agent"]), mem=capi.MEM_STAGED)
                else:
                    h.simulate_contact_network(d, hs["iv_agent"], hs["iv_cell"], hs["iv_start"], hs["iv_end"], mem=capi.MEM_HOST)
                if upload:
                    h.prefetch_intervals(nx["iv_agent"], nx["iv_cell"], nx["iv_start"], nx["iv_end"], None)
                    eng._staged = nx
                h.simulate_direct_contacts(d)
                if export:
                    rows, kv, res = eng._res[eng._k]
                    eng._k ^= 1
                    if wait_mode == "dev":          # same streams and events, destination in device memory
                        h.export_day_async(dev_rows, eng._cap_rows, dev_kv, eng._cap_exposed, dev_res, mem=capi.MEM_DEVICE)
                        continue
                    tk = h.export_day_async(rows, eng._cap_rows, kv, eng._cap_exposed, res, mem=capi.MEM_HOST)
                    cur = engine.HostDayResult(h, tk, rows, kv, res)
                    if wait_mode == "lag1":
                        if prev is not None:
                            prev.wait()
                        prev = cur
                    elif wait_mode == "now":
                        cur.wait()
        if prev is not None:
            prev.wait()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) * 1e3 / a.steps

    for upload, export, wm in ((False, False, "-"), (True, False, "-"), (False, True, "lag1"), (True, True, "lag1"),
                               (True, True, "now"), (True, True, "dev")):
        loop(upload, export, wm)
        print("upload=%-5s export=%-5s wait=%-5s  %.3f ms/day" % (upload, export, wm, loop(upload, export, wm)), flush=True)
    eng.close()


if __name__ == "__main__":
    main()
