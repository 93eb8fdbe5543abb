#!/usr/bin/env python
"""per-day stage times of the 5M-agent bench workload (one sync per day): which day streams are the long ones, and how
the K1 wall time of a day compares with the serialised sum of its launches (profiles/*launches*.csv)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mtdcovabm_b200 import capi, engine, params as prm, synth

dev = torch.device("cuda", 0)
bank = synth.load_bank()
reps = max(1, round(5_000_000 / int(bank["present"].sum())))
pop = synth.SynthPopulation(bank, reps, dev, seed=0)
streams = [pop.stream(k, 0, 1) for k in range(pop.n_days)]
eng = engine.DayEngine(pop.n_agents, pop.n_cells, max(len(s["iv_agent"]) for s in streams) + 16, prm.build_params(), device=0, seed=0)
eng.set_population(pop)
for step in range(3):
    eng.step(1 + step, streams[step % pop.n_days])
eng.h.sync()
eng.h.set_profiling(True)
for step in range(3, 3 + 2 * pop.n_days):
    eng.step(1 + step, streams[step % pop.n_days])
    ms, _ = eng.h.stage_ms()
    c = eng.h.counters()
    print("stream %d  k0 %.3f  k1 %.3f  k2 %.3f  k3 %.3f  classes %s" % (step % pop.n_days, ms["k0_csr"], ms["k1_contacts"], ms["k2_rows"],
                                                                       ms["k3_eval"], list(c.class_cells)), flush=True)
eng.close()
