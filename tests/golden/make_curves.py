"""Record the epidemic curves of the UNMODIFIED reference over 32 seeds (config 1: 1k agents + tourists, 20 days,
data/epidemiology.json) together with what is needed to re-run the contact-network + transmission + SEIR path on
the same itineraries in Philox mode (north_star: per-day two-sample KS on incidence and state occupancy).

    python tests/golden/make_curves.py [--seeds 32] [--jobs 8]      # build container only (/root/reference)
    python tests/golden/make_curves.py --name curves_10k_normal --pop 10k --variant epidemiology_normal_days.json \
           --days 30 --seeds 8 --jobs 4                              # configs 2 / 5 on the tiled 10k population

Writes tests/golden/curves_1k.npz:
  run-level : seeds, n_agents
  per seed s: s<seed>_cell_type, s<seed>_cell_indid (cell construction is seed-dependent)
              s<seed>_iv_{off,cell,agent,start,end}   interval streams of the 20 days (offsets per day)
              s<seed>_static                          age_bracket, epi_age_bracket, res_cellid, soc_rate (union over days)
              s<seed>_init_{seir,type,sev}            state at the day-1 boundary
              s<seed>_init_sched                      [(agent, day, transition, ts)] schedule entries at day 1
              s<seed>_present                         [20, N] bool: agent has a SEIR state at the boundary of day d
              s<seed>_arr_{state,sched}               arrivals: (day, agent, seir, type, sev) and their schedule entries
              s<seed>_dyn                             [20, 3]: mhd, testing_enabled, quarantine_enabled
              s<seed>_last_dose                       [(day, agent, dose_day)] changes of the last vaccination day
              s<seed>_ref_incidence [20], s<seed>_ref_occupancy [20, 6]   the reference's own curves
The itinerary stage of the reference draws from an unseeded generator (itinerary.py:43), so every run is a fresh
sample; the seed only fixes the initial state and the hot path's global RNG streams.
"""
from __future__ import annotations

import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

DAYS = 20


def one(seed, out_path, pop="1k", variant=None, days=None):
    """pop: "1k" (the reference's own population) or "10k" (make_pop10k.py tiling, configs 2 and 5)"""
    global DAYS
    import ref_harness as rh
    if days is not None:
        DAYS = days
    if pop == "10k":
        import make_pop10k
        if not os.path.exists("/tmp/pop10k/agents.json"):
            make_pop10k.tile("/tmp/pop10k", 10)
        wd = rh.prepare_workdir("/tmp/refcurves_%s_%s_%d" % (pop, (variant or "base")[:20], seed), popsubfolder="tiled10k",
                                pop_src="/tmp/pop10k", epidemiology_variant=variant)
        kw = {"popsubfolder": "tiled10k"}
    else:
        wd = rh.prepare_workdir("/tmp/refcurves_%d" % seed, epidemiology_variant=variant)
        kw = {}
    comp = {"days": []}

    def on_day(d):
        N = d["n_agents_total"]
        pre, post = d["pre"], d["post"]
        rec = dict(day=d["day"], N=N, iv_cell=d["iv_cell"].astype(np.int32), iv_agent=d["iv_agent"].astype(np.int32),
                   iv_start=d["iv_start"], iv_end=d["iv_end"], ids=pre["ids"].copy(), seir=pre["seir"].copy(),
                   inf_type=dict(pre["inf_type"]), inf_sev=dict(pre["inf_sev"]),
                   vacc={k: v[-1] for k, v in pre["vacc"].items() if len(v)},
                   pre_epi={k: v.get("state_transition_by_day") or [] for k, v in d["pre_epi"].items()},
                   static=(d["static_ids"], d["age_bracket"], d["epi_age_bracket"], d["soc_rate"], d["res_cellid"]),
                   dyn=(d["mhd"], int(d["testing_enabled"]), int(d["quarantine_enabled"])),
                   incidence=len(d["exposed_ids"]),
                   occupancy=np.bincount(post["seir"].astype(np.int64), minlength=6)[:6],
                   cell_type=d["all_cell_type"], cell_indid=d["all_cell_indid"], ext_ids=list(d["ext_ids"]),
                   ext_epi={k: v.get("state_transition_by_day") or [] for k, v in d["ext_epi"].items()})
        comp["days"].append(rec)
        # drop the heavy tape of this day: only the curves are wanted here
        d["tape_sel"] = d["tape_pair_u"] = d["tape_expo"] = None

    def on_day_drop(d):      # keep memory flat over long runs: the recorder's own copy of the day is not needed
        on_day(d)
        for k in ("pre", "post", "pre_epi", "post_epi", "ext_epi", "contacts", "iv_cell", "iv_agent"):
            d[k] = None

    rec = rh.run_reference(wd, seed=seed, days=DAYS, on_day=on_day_drop, **kw)
    rec.days.clear()
    days = comp["days"]
    assert len(days) == DAYS
    N = days[0]["N"]
    out = {}
    ct = days[0]["cell_type"]; nc = max(ct) + 1
    cell_type = np.full(nc, -1, np.int8); cell_indid = np.full(nc, -1, np.int16)
    for k, v in ct.items():
        cell_type[k] = v
    for k, v in days[0]["cell_indid"].items():
        cell_indid[k] = v
    out["cell_type"], out["cell_indid"] = cell_type, cell_indid
    off = np.zeros(DAYS + 1, np.int64)
    for i, d in enumerate(days):
        off[i + 1] = off[i] + len(d["iv_cell"])
    out["iv_off"] = off
    small = nc < 32768 and N < 32768
    out["iv_cell"] = np.concatenate([d["iv_cell"] for d in days]).astype(np.int16 if small else np.int32)
    out["iv_agent"] = np.concatenate([d["iv_agent"] for d in days]).astype(np.int16 if small else np.int32)
    out["iv_start"] = np.concatenate([d["iv_start"] for d in days]); out["iv_end"] = np.concatenate([d["iv_end"] for d in days])
    age = np.zeros(N, np.int8); epi = np.zeros(N, np.int8); soc = np.ones(N, np.float64); res = np.full(N, -1, np.int32)
    for d in days:
        ids, a, e, s, r = d["static"]
        age[ids] = a; epi[ids] = e; soc[ids] = s; res[ids] = r
    out["age_bracket"], out["epi_age_bracket"], out["soc_rate"], out["res_cellid"] = age, epi, soc, res
    present = np.zeros((DAYS, N), bool)
    arr_state, arr_sched, dose = [], [], []
    last = {}
    for i, d in enumerate(days):
        present[i, d["ids"]] = True
        new = d["ids"] if i == 0 else d["ids"][~present[i - 1, d["ids"]]]
        seir = dict(zip(d["ids"].tolist(), d["seir"].tolist()))
        # arrivals + agents whose state the itinerary stage set outside the hot path (tourist entry infection,
        # re-entry infection of returning locals, day-1 infectious: itinerary.py:392,645,1578,1584)
        inject = sorted(set(new.tolist()) | set(a for a in d["ext_ids"] if a in seir))
        for a in inject:
            arr_state.append((d["day"], a, seir[a], d["inf_type"].get(a, 0), d["inf_sev"].get(a, 0)))
            ent = d["ext_epi"].get(a) if a in d["ext_epi"] else d["pre_epi"].get(a, [])
            for (dd, t, ts) in ent or []:
                arr_sched.append((d["day"], a, dd, t, ts))
        for a, v in d["vacc"].items():
            if last.get(a) != v:
                dose.append((d["day"], a, v)); last[a] = v
    out["present"] = np.packbits(present, axis=1)
    out["arr_state"] = np.array(arr_state, np.int32).reshape(-1, 5)
    out["arr_sched"] = np.array(arr_sched, np.int32).reshape(-1, 5)
    out["last_dose"] = np.array(dose, np.int32).reshape(-1, 3)
    out["dyn"] = np.array([d["dyn"] for d in days], np.float64)
    out["days"] = np.array([DAYS], np.int64)
    out["ref_incidence"] = np.array([d["incidence"] for d in days], np.int32)
    out["ref_occupancy"] = np.stack([d["occupancy"] for d in days]).astype(np.int32)
    out["n_agents"] = np.array([N], np.int64)
    np.savez_compressed(out_path, **out)


def main():
    def opt(name, default=None):
        return sys.argv[sys.argv.index(name) + 1] if name in sys.argv else default

    pop, variant, days = opt("--pop", "1k"), opt("--variant"), opt("--days")
    days = None if days is None else int(days)
    if len(sys.argv) > 1 and sys.argv[1] == "--one":
        one(int(sys.argv[2]), sys.argv[3], pop, variant, days)
        return
    n_seeds = int(opt("--seeds", 32))
    jobs = int(opt("--jobs", 8))
    seeds = list(range(100, 100 + n_seeds))
    name = opt("--name", "curves_1k")
    tmp = "/tmp/refcurves_out_" + name
    os.makedirs(tmp, exist_ok=True)
    procs = []
    pending = list(seeds)
    while pending or procs:
        while pending and len(procs) < jobs:
            s = pending.pop(0)
            extra = ["--pop", pop] + (["--variant", variant] if variant else []) + (["--days", str(days)] if days else [])
            procs.append((s, subprocess.Popen([sys.executable, os.path.abspath(__file__), "--one", str(s),
                                               os.path.join(tmp, "s%d.npz" % s)] + extra, stdout=subprocess.DEVNULL)))
        s, p = procs.pop(0)
        if p.wait() != 0:
            raise SystemExit("seed %d failed" % s)
    merged = {"seeds": np.array(seeds, np.int32)}
    for s in seeds:
        z = np.load(os.path.join(tmp, "s%d.npz" % s))
        for k in z.files:
            if k in ("n_agents", "days"):
                if k in merged:
                    assert np.array_equal(merged[k], z[k])
                merged[k] = z[k]
            else:
                merged["s%d_%s" % (s, k)] = z[k]
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **merged)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()
