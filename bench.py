#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 contact-network + transmission + SEIR day step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--agents A]

One "step" = one simulated day of the hot path over a synthetic population (K4 SEIR advance, K0 CSR build,
K1 per-cell contact network, K2 rows, K3 transmission + deltas [+ all-gather + apply when N > 1]).
Workload (BASELINE.json configs[3], the configuration the metric's target is quoted on): 5M active agents =
the reference's own 1k-agent itinerary streams tiled 5000x (mtdcovabm_b200/synth.py), cells sharded over the N
ranks (strong scaling: the population is fixed, each rank owns 1/N of the cells), state replicated.

Prints ONE JSON line (rank 0).  `value` = agent-contact pairs evaluated per second, whole job, inputs resident in
HBM; `e2e` = the same through the public engine API with HOST buffers (H2D of the day's intervals and D2H of the
contact rows + exposed ids inside the timed region).  `--impl reference` times the CPU oracle port of the reference
path (the reference is pure Python and its sources do not travel to the GPU box) on all host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import mtdcovabm_b200  # noqa: E402,F401  (sets CUDA_DEVICE_MAX_CONNECTIONS for multi-rank runs before CUDA is initialised)

METRIC = "agent_contact_pairs_per_s"
UNIT = "pairs/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--agents", type=int, default=5_000_000, help="active agents of the synthetic population")
    ap.add_argument("--cpu-sample-agents", type=int, default=0, help="agents of the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--shard-of", type=int, default=0,
                    help="development aid: run rank 0's shard of a W-rank job on ONE GPU (no exchange); the line is not a bench value")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        try:
            for line in self.proc.stdout:
                f = [x.strip() for x in line.strip().split(",")]
                if len(f) >= 7:
                    self.rows.append((time.perf_counter(), f))
        except Exception:
            pass

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t.start()
            time.sleep(0.3)              # let the first samples arrive before the timed region starts
        except Exception:
            self.proc = None
        self.t0 = time.perf_counter()
        return self

    def __exit__(self, *a):
        self.t1 = time.perf_counter()
        if self.proc is not None:
            time.sleep(0.06)
            self.proc.terminate()
            self.t.join(timeout=3)

    def summary(self):
        rows = [f for (t, f) in self.rows if self.t0 <= t <= self.t1 + 0.06] or [f for (_, f) in self.rows[-1:]]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons,
                "power_w_max": max(float(r[2]) for r in rows), "samples": len(rows)}


# ------------------------------------------------------------------------------------------- CPU arm
def cpu_day_numpy(bank, reps, k, seed):
    """bank day k tiled `reps` times on the host (same construction as synth.SynthPopulation, numpy)"""
    N1 = int(bank["n_agents"][0]); C1 = len(bank["cell_type"])
    a1 = bank["d%d_iv_agent" % k].astype(np.int64); c1 = bank["d%d_iv_cell" % k].astype(np.int64)
    M1 = len(a1)
    r = np.repeat(np.arange(reps, dtype=np.int64), M1)
    agent = np.tile(a1, reps) + r * N1
    cell = (np.tile(c1, reps) + r * C1).astype(np.int32)
    start = np.tile(bank["d%d_iv_start" % k], reps); end = np.tile(bank["d%d_iv_end" % k], reps)
    rng = np.random.default_rng(1234 + seed)
    perm = rng.permutation(N1 * reps)
    agent = perm[agent]
    o = np.argsort(agent, kind="stable")
    return (agent[o].astype(np.int32), cell[o], start[o], end[o], perm)


def cpu_oracle_run(args, steps, warmup, sample_agents):
    """Times the CPU oracle port (oracle/mtd_oracle.c: contact network + transmission + SEIR advance) with one
    worker thread per host core, cells partitioned over the workers like contactnetwork_mp.contactnetwork_parallel
    (contactnetwork_mp.py:59-190).  -> dict(value pairs/s, cores, sample, ms_per_step)"""
    from concurrent.futures import ThreadPoolExecutor
    from mtdcovabm_b200 import params as prm, synth
    from oracle import mtd_oracle as mo
    bank = synth.load_bank()
    P = prm.build_params()
    N1 = int(bank["n_agents"][0])
    n_present1 = int(bank["present"].sum())
    reps = max(1, sample_agents // n_present1)
    cores = os.cpu_count() or 1
    nb = int(bank["n_days"])
    days = []
    for k in range(nb):
        agent, cell, start, end, perm = cpu_day_numpy(bank, reps, k, args.seed)
        part = (cell.astype(np.int64) * 2654435761 & 0xFFFFFFFF) >> 7
        shards = []
        for w in range(cores):
            sel = (part % cores) == w
            shards.append(tuple(np.ascontiguousarray(x[sel]) for x in (cell, agent, start, end)))
        days.append(shards)
    N = N1 * reps
    st = mo.AgentState(N)
    for name in ("age_bracket", "epi_age_bracket", "soc_rate"):
        getattr(st, name)[perm] = np.tile(bank[name], reps)
    present = np.zeros(N, bool); present[perm] = np.tile(bank["present"], reps)
    rng = np.random.default_rng(99 + args.seed)
    st.seir[:] = np.where(present, 1, 0)
    inf = present & (rng.random(N) < 0.01)
    st.seir[inf] = 3; st.inf_type[inf] = 3; st.inf_sev[inf] = 1
    cell_type = np.tile(bank["cell_type"], reps); cell_indid = np.tile(bank["cell_indid"], reps)
    mo.lib()

    def worker(shard, day):
        cell, agent, start, end = shard
        ctx = mo.make_ctx(mo.MODE_PHILOX, day, 0.0, False, False, seed=args.seed)
        rows, dur, rc, s1 = mo.contact_network(P, ctx, cell, agent, start, end, cell_type, cell_indid, st)
        return rows, dur, rc, s1.pairs_evaluated, shard

    total_pairs, t_total = 0, 0.0
    with ThreadPoolExecutor(cores) as ex:
        for step in range(warmup + steps):
            day = 1 + step
            shards = days[step % nb]
            t0 = time.perf_counter()
            mo.advance(day, st, None)
            res = list(ex.map(lambda sh: worker(sh, day), shards))
            # transmission mutates shared per-agent state: applied in partition order on one thread, like the
            # reference's merge of the workers' results (daskutil.py:168)
            ctx = mo.make_ctx(mo.MODE_PHILOX, day, 0.0, False, False, seed=args.seed)
            for rows, dur, rc, _, (cell, agent, start, end) in res:
                mo.transmission(P, ctx, cell, agent, start, end, cell_type, rows, dur, rc, st)
            dt = time.perf_counter() - t0
            if step >= warmup:
                t_total += dt
                total_pairs += sum(r[3] for r in res)
    return dict(value=total_pairs / t_total, cores=cores, ms_per_step=1e3 * t_total / steps,
                sample="%d active agents (%d replicas of the 1k itinerary bank), %d day-steps" % (n_present1 * reps, reps, steps),
                agents=n_present1 * reps)


def parity_on_sample(args, sample_agents):
    """The oracle (one thread, the reference's cell order) and the CUDA path on the SAME shard of the bench workload: day
    stream 0 of a `sample_agents` tiling from the same initial state, Philox mode.  Contact rows must be equal element for
    element and in order, exposed ids likewise.  -> dict for the JSON line; raises on a mismatch."""
    from mtdcovabm_b200 import capi, params as prm, synth
    from oracle import mtd_oracle as mo
    bank = synth.load_bank()
    P = prm.build_params()
    N1 = int(bank["n_agents"][0]); n_present1 = int(bank["present"].sum())
    reps = max(1, sample_agents // n_present1)
    agent, cell, start, end, perm = cpu_day_numpy(bank, reps, 0, args.seed)
    N = N1 * reps
    st = mo.AgentState(N)
    for name in ("age_bracket", "epi_age_bracket", "soc_rate"):
        getattr(st, name)[perm] = np.tile(bank[name], reps)
    rc = np.tile(bank["res_cellid"].astype(np.int64), reps)
    rc = np.where(rc >= 0, rc + np.repeat(np.arange(reps, dtype=np.int64), N1) * len(bank["cell_type"]), rc)
    st.res_cellid[perm] = rc.astype(np.int32)
    present = np.zeros(N, bool); present[perm] = np.tile(bank["present"], reps)
    rng = np.random.default_rng(99 + args.seed)
    st.seir[:] = np.where(present, 1, 0)
    inf = present & (rng.random(N) < 0.02)
    st.seir[inf] = 3; st.inf_type[inf] = 3; st.inf_sev[inf] = 1
    cell_type = np.ascontiguousarray(np.tile(bank["cell_type"], reps), np.int8)
    cell_indid = np.ascontiguousarray(np.tile(bank["cell_indid"], reps), np.int16)
    h = capi.Handle(N, len(cell_type), len(agent) + 16)
    h.set_params(P)
    h.set_cells(cell_type, cell_indid)
    h.set_static(st.age_bracket, st.epi_age_bracket, st.soc_rate, st.res_cellid)
    h.set_state(seir=st.seir, inf_type=st.inf_type, inf_sev=st.inf_sev)
    d = h.make_day(capi.MODE_PHILOX, 1, 0.0, False, False, seed=args.seed)
    h.simulate_contact_network(d, agent, cell, start, end)
    h.simulate_direct_contacts(d)
    c = h.counters()
    grow, gdur, gcell = h.get_contacts()
    gexp = h.get_exposed()
    h.close()
    ctx = mo.make_ctx(mo.MODE_PHILOX, 1, 0.0, False, False, seed=args.seed)
    orow, odur, orc, s1 = mo.contact_network(P, ctx, cell, agent, start, end, cell_type, cell_indid, st)
    oexp, _, _ = mo.transmission(P, ctx, cell, agent, start, end, cell_type, orow, odur, orc, st)
    ok_rows = bool(np.array_equal(grow, orow) and np.array_equal(gdur, odur) and np.array_equal(gcell, orc))
    ok_exp = bool(np.array_equal(gexp, oexp))
    ok_cnt = bool(int(c.pairs_evaluated) == int(s1.pairs_evaluated) and int(c.n_potential) == int(s1.n_potential))
    out = {"sample": "%d active agents (day stream 0 of the bench workload, %d replicas)" % (n_present1 * reps, reps),
           "rows": int(len(orow)), "exposed": int(len(oexp)), "pairs": int(s1.pairs_evaluated), "rows_equal": ok_rows,
           "exposed_equal": ok_exp, "counters_equal": ok_cnt}
    if not (ok_rows and ok_exp and ok_cnt):
        raise SystemExit("bench.py: the CUDA path and the oracle disagree on the bench workload sample: %s" % json.dumps(out))
    return out


def reference_python_timing():
    """the UNMODIFIED reference's own contactnetwork_mp.contactnetwork_parallel, timed by tools/time_reference.py in the
    build container (its sources do not travel to the GPU box): carried here next to the C port's number"""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r02_reference_python_timing.json")))
    except Exception:
        return None


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample = args.cpu_sample_agents or 1_000_000
    r = cpu_oracle_run(args, max(1, args.steps), max(1, min(args.warmup, 2)), sample)
    # same `config` as the B200 arm (the workload the metric is quoted on); the bounded sample each step really ran is
    # described in cpu_baseline.sample
    from mtdcovabm_b200 import synth
    bank = synth.load_bank()
    n1 = int(bank["present"].sum())
    reps_full = max(1, round(args.agents / n1))
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "int32+f64", "data": "synthetic",
            "config": workload_config(args, n1 * reps_full, int(len(bank["d0_iv_agent"])) * reps_full,
                                      int(len(bank["cell_type"])) * reps_full),
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                             "sample": r["sample"], "reference_python": reference_python_timing()},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(args, agents, intervals, cells):
    par = "cells sharded x%d, state replicated" % args.gpus
    return {"workload": "synthetic %d-agent population: reference 1k-agent itinerary bank tiled, ids shuffled, "
                        "agent-major emission order (BASELINE.json configs[3])" % agents,
            "agents": agents, "intervals_per_day": intervals, "cells": cells, "parallelism": par,
            "rng": "philox4x32-10", "l2": "inputs larger than L2 (7 distinct day streams cycled)"}


# ------------------------------------------------------------------------------------------- GPU arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    from mtdcovabm_b200 import capi, engine, params as prm, synth

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the B200 path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    bank = synth.load_bank()
    n_present1 = int(bank["present"].sum())
    reps = max(1, round(args.agents / n_present1))
    pop = synth.SynthPopulation(bank, reps, dev, seed=args.seed)
    nb = pop.n_days
    streams = [pop.stream(k, rank, world) if not (args.shard_of and world == 1) else pop.stream(k, 0, args.shard_of) for k in range(nb)]
    max_m = max(len(s["iv_agent"]) for s in streams)
    P = prm.build_params()
    eng = engine.DayEngine(pop.n_agents, pop.n_cells, max_m + 16, P, device=local, rank=rank, world=world,
                           seed=args.seed)
    eng.set_population(pop)
    h = eng.h

    # calibration: the pairs a day's stream holds do not depend on the epidemic state -> count them once
    pairs = []
    d = h.make_day(capi.MODE_PHILOX, 1, 0.0, False, False, seed=args.seed)
    for k in range(nb):
        s = streams[k]
        with torch.cuda.stream(eng.stream):
            h.simulate_contact_network(d, s["iv_agent"], s["iv_cell"], s["iv_start"], s["iv_end"], iv_order=s["iv_order"],
                                       mem=capi.MEM_DEVICE)
        c = h.counters()
        pairs.append((int(c.pairs_evaluated), int(c.n_rows), int(c.n_slots), int(c.n_incidences)))
    eng.set_population(pop)          # reset the state
    torch.cuda.synchronize()
    eng.set_inputs_ready(True)       # the day streams were built above and are complete: K0/K1 of the coming days may overlap the current one

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---------------- device-resident timed loop
    for step in range(args.warmup):
        eng.step(1 + step, streams[step % nb])
    barrier()
    h.set_profiling(True)
    stage_sum = {}
    launches = 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        ev0.record(eng.stream)
        t_sub0 = time.perf_counter()
        for step in range(args.warmup, args.warmup + args.steps):
            eng.step(1 + step, streams[step % nb])
        host_submit_ms = (time.perf_counter() - t_sub0) * 1e3 / args.steps   # host time to enqueue one day (no waiting in it unless a queue is full)
        ev1.record(eng.stream)
        barrier()
    ms = allmax(ev0.elapsed_time(ev1))
    local_pairs = sum(pairs[step % nb][0] for step in range(args.warmup, args.warmup + args.steps))
    total_pairs = allsum(local_pairs)
    value = total_pairs / (ms * 1e-3)
    # per-stage device time, mean over the timed steps (CUDA events on the engine's stream, recorded by the library
    # around every stage), and the kernels launched over the loop
    st_ms, st_launch = h.stage_ms()
    launches = sum(st_launch.values())
    c = h.counters()

    # roofline of the dominant kernel group (K1): algorithmic bytes = 10*M + 13*U + 12*D  (SURVEY 8d),
    # mean over the timed steps (this rank's shard)
    timed = range(args.warmup, args.warmup + args.steps)
    M_mean = sum(len(streams[t % nb]["iv_agent"]) for t in timed) / args.steps
    D_mean = sum(pairs[t % nb][1] for t in timed) / args.steps
    P_mean = sum(pairs[t % nb][0] for t in timed) / args.steps
    U_mean = sum(pairs[t % nb][3] for t in timed) / args.steps        # (cell, agent) incidences, counted by K1
    k1_bytes = 10 * M_mean + 13 * U_mean + 12 * D_mean
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # measured DRAM traffic of the K1 launches of one day (ncu --set full capture, committed under profiles/), scaled to
    # this rank's share of the day's intervals
    traffic = None
    warp_instr = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "r02_k1_traffic.json")))
        traffic = tr["dram_bytes_total"] * M_mean / tr["intervals_of_day"]
        warp_instr = tr["warp_instructions"] * M_mean / tr["intervals_of_day"]
    except Exception:
        pass
    # K1 stages of consecutive days overlap in the timed region (the chain kernels of day d run beside the prepare and
    # small-cell kernels of day d+1), so the span of one stage -- its first kernel to its last -- can exceed the period of the
    # loop.  The time one day's K1 costs the loop is at most the period: the roofline uses min(span, period).
    k1_span_ms = st_ms["k1_contacts"]
    k1_ms = min(k1_span_ms, ms / args.steps) if k1_span_ms > 0 else 0.0
    achieved = k1_bytes / (k1_ms * 1e-3) / 1e9 if k1_ms > 0 else 0.0
    roofline = {"bound": "hbm", "kernel": "k1_contacts (3 sub-warp kernels for cells <= 32 intervals; per class of 33..320 intervals a block-per-cell prepare kernel and an 8-lanes-per-cell chain kernel; 2 block-per-cell kernels above; 6 ordering helpers)", "achieved": achieved,
                "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": "ncu dram__bytes_read.sum + dram__bytes_write.sum over the K1 launches, mean per day over whole days of the bench (profiles/r02_k1_traffic.json, launch list profiles/r02_k1_traffic_launches.csv), scaled by this rank's intervals",
                "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback",
                "algorithmic_bytes_per_day": k1_bytes, "algorithmic_units": {"M_intervals": M_mean, "U_incidences": U_mean, "D_rows": D_mean},
                "k1_ms": k1_ms, "k1_span_ms": k1_span_ms, "stage_ms": st_ms,
                "k1_pairs_per_s": P_mean / (k1_ms * 1e-3) if k1_ms > 0 else None,
                "note": "K1 is issue/latency-bound (sequential greedy sampling per cell), not HBM-bound: see DESIGN.md"}
    # SM-side bound of K1: warp instructions issued per day (ncu smsp__inst_executed.sum, same capture as
    # `traffic`) against the issue rate of the device, 4 schedulers x 148 SMs x the SM clock sampled during the loop
    cs = clocks.summary()
    if warp_instr and k1_ms > 0 and cs.get("sm_mhz"):
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        issue_peak = sms * 4 * cs["sm_mhz"] * 1e6
        roofline["sm_side"] = {"bound": "issue", "warp_instructions_per_day": warp_instr,
                               "achieved_warp_inst_per_s": warp_instr / (k1_ms * 1e-3), "peak_warp_inst_per_s": issue_peak,
                               "frac": warp_instr / (k1_ms * 1e-3) / issue_peak,
                               "warp_instructions_per_pair": warp_instr / P_mean if P_mean else None,
                               "note": "k1_ms = min(span of the stage, period of the loop): K1's launches run on about ten concurrent streams and overlap the next day's"}
    # the timed loop runs K0 and K1 of the coming days beside K2/K3 of the current one, so its per-stage times include the sharing; a short
    # untimed pass with the stages one after the other gives the stand-alone stage times next to them
    eng.set_inputs_ready(False)
    for step in range(args.warmup + args.steps, args.warmup + args.steps + 2 * nb):
        eng.step(1 + step, streams[step % nb])
    ser_ms, _ = h.stage_ms()
    eng.set_inputs_ready(True)
    roofline["stage_ms_serial"] = ser_ms
    if ser_ms.get("k1_contacts", 0) > 0:      # the same algorithmic bytes over K1's stand-alone time, next to the pipelined one
        roofline["k1_ms_serial"] = ser_ms["k1_contacts"]
        roofline["achieved_serial"] = k1_bytes / (ser_ms["k1_contacts"] * 1e-3) / 1e9
        roofline["frac_serial"] = roofline["achieved_serial"] / peak
    roofline["stage_note"] = ("stage_ms / k1_span_ms: CUDA events around each stage over the timed region, where the day runs as a "
                              "pipeline (K0 of day d+2, K1 of days d+1 and d and K2/K3 of day d share the SMs; the state "
                              "chain k3_apply / k4_advance has its own stream), so a stage's time includes waiting for SM slots and, for K1, "
                              "the overlap with the next day's stage; k1_ms = min(k1_span_ms, ms_per_step); "
                              "stage_ms_serial: the same events over %d untimed days with the stages one after the other" % (2 * nb))
    h.set_profiling(False)

    # ---------------- end to end: host buffers in, host results out, every step
    e2e = None
    if not args.no_e2e:
        host = []
        for s in streams:
            hs = {k: (v.cpu().pin_memory() if v is not None else None) for k, v in s.items() if k != "n_total"}
            host.append(hs)
        cap_rows = max(p[1] for p in pairs) + 1024
        h2d = d2h = 0
        eng.set_population(pop)
        # pipelined host boundary: day d+1's intervals are uploaded and day d-1's results land in pinned host
        # memory while day d computes; every result is waited for and read inside the timed region
        eng.enable_host_pipeline(cap_rows)
        api = "engine.DayEngine.step_host (pinned host intervals in; packed rows, exposed ids, occupancy out to pinned host memory)"

        def run(first, count, account):
            nonlocal h2d, d2h
            prev = None
            for step in range(first, first + count):
                s = host[step % nb]
                res = eng.step_host(1 + step, s, next_s=host[(step + 1) % nb])
                if prev is not None:
                    prev.wait()
                    if account:
                        d2h += prev.nbytes
                if account:
                    h2d += sum(v.numel() * v.element_size() for v in s.values() if v is not None)
                prev = res
            prev.wait()
            if account:
                d2h += prev.nbytes
            return prev

        run(0, min(args.warmup, 3), False)
        barrier()
        t0 = time.perf_counter()
        last_res = run(args.warmup, args.steps, True)
        barrier()
        e_ms = allmax((time.perf_counter() - t0) * 1e3)
        assert last_res.n_rows > 0 and int(last_res.occupancy.sum()) > 0
        e2e = {"value": total_pairs / (e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(allsum(h2d)) // args.steps,
               "d2h_bytes_per_step": int(allsum(d2h)) // args.steps, "ms_per_step": e_ms / args.steps, "api": api,
               "mode": "pipelined: day d+1's intervals are uploaded while day d computes (needs tomorrow's itinerary today: "
                       "legal for recorded / tiled streams, not behind sim.py's loop)"}
        # the dependency sim.py's day loop really has (the itinerary of day d+1 reads the state after day d,
        # itinerary.py:2283-2528): no prefetch, the result of day d is waited for before day d+1 is submitted
        ser_steps = max(2, min(args.steps, 2 * nb))

        def run_serial(first, count):
            for step in range(first, first + count):
                r = eng.step_host(1 + step, host[step % nb], next_s=None)
                r.wait()
            return r

        run_serial(args.warmup + args.steps, 2)
        barrier()
        t0 = time.perf_counter()
        run_serial(args.warmup + args.steps + 2, ser_steps)
        barrier()
        s_ms = allmax((time.perf_counter() - t0) * 1e3)
        ser_pairs = allsum(sum(pairs[st_ % nb][0] for st_ in range(args.warmup + args.steps + 2, args.warmup + args.steps + 2 + ser_steps)))
        # the same pipelined loop for a caller that keeps the contact rows in HBM (contact tracing on the device):
        # counts, occupancy and exposed ids only
        eng.enable_host_pipeline(cap_rows, export_rows=False)
        d2h_keep = d2h
        run(args.warmup + args.steps + 2 + ser_steps, 2, False)
        barrier()
        t0 = time.perf_counter()
        run(args.warmup + args.steps + 4 + ser_steps, ser_steps, False)
        barrier()
        n_ms = allmax((time.perf_counter() - t0) * 1e3)
        d2h = d2h_keep
        nr_pairs = allsum(sum(pairs[st_ % nb][0] for st_ in range(args.warmup + args.steps + 4 + ser_steps, args.warmup + args.steps + 4 + 2 * ser_steps)))
        e2e["rows_on_device"] = {"value": nr_pairs / (n_ms * 1e-3), "unit": UNIT, "ms_per_step": n_ms / ser_steps, "steps": ser_steps,
                                 "mode": "pipelined, contact rows stay in HBM for device-side contact tracing: intervals in; counts, occupancy, exposed ids out"}
        e2e["serial"] = {"value": ser_pairs / (s_ms * 1e-3), "unit": UNIT, "ms_per_step": s_ms / ser_steps, "steps": ser_steps,
                         "mode": "dependency-honouring: upload, compute and export of day d complete before day d+1 is submitted"}

    cpu = None
    parity = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_oracle_run(args, 20, 2, args.cpu_sample_agents or 1_000_000)     # ~10 s of CPU work
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"],
               "reference_python": reference_python_timing()}
        parity = parity_on_sample(args, min(args.cpu_sample_agents or 1_000_000, 1_000_000))
    # per-rank stage times of the timed region: load balance of K1 and the wait for the slowest peer inside k3_apply
    per_rank = None
    if world > 1:
        allst = [None] * world
        dist.all_gather_object(allst, {k: st_ms.get(k) for k in ("k0_csr", "k1_contacts", "k2_rows", "k3_eval", "k3_apply", "k4_advance")})
        per_rank = {k: [a.get(k) for a in allst] for k in allst[0]}

    contacts_per_day = allsum(allsum_rows(pairs, args, nb))
    if rank == 0:
        agents = pop.n_present
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "int32+f64", "data": "synthetic",
                "config": workload_config(args, agents, int(streams[0]["n_total"]), pop.n_cells),
                "exchange": {"none": "single GPU", "p2p": "exposure records pushed with stores into NVLink peer memory (CUDA IPC), arrival flags",
                             "nccl": "NCCL all-gather of fixed-capacity record buffers"}[eng.exchange],
                "agent_days_per_s": agents * args.steps / (ms * 1e-3),
                "pairs_per_day": total_pairs / args.steps, "contacts_per_day": contacts_per_day,
                "exposed_last_day": int(c.n_exposed),
                "roofline": roofline, "cpu_baseline": cpu, "parity": parity, "per_rank_stage_ms": per_rank, "e2e": e2e,
                "gpu_launches": launches, "host_submit_ms_per_step": host_submit_ms,
                "clocks": clocks.summary()}
        print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def allsum_rows(pairs, args, nb):
    return sum(pairs[step % nb][1] for step in range(args.warmup, args.warmup + args.steps)) / args.steps


if __name__ == "__main__":
    # the JSON line is the only thing on stdout: native libraries (NCCL's version banner) write to fd 1 too, so fd 1 is
    # pointed at stderr for the run and the line goes to the saved descriptor
    sys.stdout.flush()
    _real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = _real_stdout
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
