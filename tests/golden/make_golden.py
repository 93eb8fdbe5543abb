"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference in this container.

Usage (build container only; /root/reference must exist):
    python tests/golden/make_golden.py            # config 1 (seed 6, 20 days) + strict variant (seed 7, 25 days)

Each run writes  tests/golden/<name>.npz  holding, per recorded day `dNN_*`:
  inputs at the ContactNetwork.simulate_contact_network boundary (contactnetwork.py:65), the reference's outputs
  (contact rows, exposed ids, post state / per-agent records) and the keyed draw tape;
plus run-level arrays (cell types, static agent attributes) and the update_agent_state KAT (seirstateutil.py:38).
The file format is decoded by tests/golden/golden_io.py.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness as rh  # noqa: E402

NONE = -32768  # "None" marker for day-valued fields


def enc_epi(epi, ids):
    """Per-agent records -> flat arrays.  ids: sorted active agent ids."""
    n = len(ids)
    sch_off = np.zeros(n + 1, dtype=np.int32)
    sch = []
    test = np.full((n, 2), NONE, dtype=np.int32)
    tres = np.full((n, 2), NONE, dtype=np.int32)
    quar = np.full((n, 3), NONE, dtype=np.int32)
    hosp = np.full((n, 3), NONE, dtype=np.int32)
    present = np.zeros(n, dtype=np.int8)
    for i, a in enumerate(ids):
        e = epi.get(int(a))
        if e is not None:
            present[i] = 1
            for ent in (e.get("state_transition_by_day") or []):
                sch.append(ent)
            for arr, key in ((test, "test_day"), (tres, "test_result_day"), (quar, "quarantine_days"),
                             (hosp, "hospitalisation_days")):
                v = e.get(key)
                if v is not None and len(v) > 0:
                    arr[i, :len(v)] = v
        sch_off[i + 1] = len(sch)
    sch = np.array(sch, dtype=np.int32).reshape(-1, 3)
    return {"epi_present": present, "sch_off": sch_off, "sch": sch, "test": test, "tres": tres, "quar": quar,
            "hosp": hosp}


def enc_state(st, N):
    seir = np.zeros(N, dtype=np.int8)
    seir[st["ids"]] = st["seir"]
    it = np.zeros(N, dtype=np.int8)
    sv = np.zeros(N, dtype=np.int8)
    itp = np.zeros(N, dtype=np.int8)
    svp = np.zeros(N, dtype=np.int8)
    for k, v in st["inf_type"].items():
        it[k] = v; itp[k] = 1
    for k, v in st["inf_sev"].items():
        sv[k] = v; svp[k] = 1
    ld = np.full(N, NONE, dtype=np.int32)
    for k, v in st["vacc"].items():
        if len(v) > 0:
            ld[k] = v[-1]
    return {"seir": seir, "inf_type": it, "inf_sev": sv, "inf_type_present": itp, "inf_sev_present": svp,
            "last_dose": ld}


def encode_day(d, out, prefix):
    N = d["n_agents_total"]
    for k in ("iv_cell", "iv_agent", "iv_start", "iv_end", "contacts", "exposed_ids"):
        out[prefix + k] = d[k]
    out[prefix + "scalars"] = np.array([d["day"], d["weekday"], int(d["testing_enabled"]),
                                        int(d["quarantine_enabled"]), int(d["contact_tracing_enabled"]), N,
                                        d["n_locals"]], dtype=np.int64)
    out[prefix + "mhd"] = np.array([d["mhd"]], dtype=np.float64)
    ids = d["static_ids"]
    out[prefix + "act_ids"] = ids
    out[prefix + "age_bracket"] = d["age_bracket"]
    out[prefix + "epi_age_bracket"] = d["epi_age_bracket"]
    out[prefix + "soc_rate"] = d["soc_rate"]
    out[prefix + "res_cellid"] = d["res_cellid"]
    for tag, st, epi in (("pre_", d["pre"], d["pre_epi"]), ("post_", d["post"], d["post_epi"])):
        for k, v in enc_state(st, N).items():
            out[prefix + tag + k] = v
        for k, v in enc_epi(epi, ids).items():
            out[prefix + tag + k] = v
    # tape: selections
    sel = d["tape_sel"]
    out[prefix + "sel_cell"] = np.array([s[0] for s in sel], dtype=np.int32)
    out[prefix + "sel_agent"] = np.array([s[1] for s in sel], dtype=np.int32)
    out[prefix + "sel_ncand"] = np.array([s[2] for s in sel], dtype=np.int32)
    off = np.zeros(len(sel) + 1, dtype=np.int32)
    picks = []
    for i, s in enumerate(sel):
        picks.extend(s[3]); off[i + 1] = len(picks)
    out[prefix + "sel_off"] = off
    out[prefix + "sel_picks"] = np.array(picks, dtype=np.int32)
    pu = d["tape_pair_u"]
    out[prefix + "pair_key"] = np.array([[p[0], p[1], p[2]] for p in pu], dtype=np.int32).reshape(-1, 3)
    out[prefix + "pair_u"] = np.array([p[3] for p in pu], dtype=np.float64)
    ex = d["tape_expo"]
    out[prefix + "expo_key"] = np.array([[e[0], e[1], e[2], e[3]] for e in ex], dtype=np.int32).reshape(-1, 4)
    out[prefix + "expo_row"] = np.array([e[4] for e in ex], dtype=np.float64).reshape(-1, rh.EXPO_W)
    eo = np.zeros(len(ex) + 1, dtype=np.int32)
    ets = []
    for i, e in enumerate(ex):
        ets.extend(e[5]); eo[i + 1] = len(ets)
    out[prefix + "expo_ts_off"] = eo
    out[prefix + "expo_ts"] = np.array(ets, dtype=np.int16)


def encode_run(rec, keep_days):
    out = {}
    d0 = rec.days[0]
    ct = d0["all_cell_type"]
    ncell = max(ct.keys()) + 1
    cell_type = np.full(ncell, -1, dtype=np.int8)
    for k, v in ct.items():
        cell_type[k] = v
    cell_indid = np.full(ncell, -1, dtype=np.int16)
    for k, v in d0["all_cell_indid"].items():
        cell_indid[k] = v
    out["cell_type"] = cell_type
    out["cell_indid"] = cell_indid
    days = [d for d in rec.days if d["day"] in keep_days]
    out["days"] = np.array([d["day"] for d in days], dtype=np.int32)
    for d in days:
        encode_day(d, out, "d%02d_" % d["day"])
    # update_agent_state KAT
    adv = rec.advance_calls
    if len(adv) > 20000:        # long runs: the KAT keeps the calls of the kept days only
        ks = set(keep_days)
        adv = [a for a in adv if a[0] in ks]
    A = len(adv)
    out["adv_day_agent"] = np.array([[a[0], a[1]] for a in adv], dtype=np.int32).reshape(-1, 2)
    out["adv_pre"] = np.array([a[3] for a in adv], dtype=np.int8).reshape(-1, 3)
    out["adv_post"] = np.array([a[4] for a in adv], dtype=np.int8).reshape(-1, 3)
    out["adv_applied"] = np.array([int(a[6]) for a in adv], dtype=np.int8)
    bo = np.zeros(A + 1, dtype=np.int32); ao = np.zeros(A + 1, dtype=np.int32)
    b, af = [], []
    for i, a in enumerate(adv):
        b.extend(a[2]); af.extend(a[5]); bo[i + 1] = len(b); ao[i + 1] = len(af)
    out["adv_before_off"] = bo; out["adv_before"] = np.array(b, dtype=np.int32).reshape(-1, 3)
    out["adv_after_off"] = ao; out["adv_after"] = np.array(af, dtype=np.int32).reshape(-1, 3)
    if rec.traces:
        encode_traces(rec, out)
    return out


def encode_traces(rec, out, days_back=2):
    """contact_tracing calls (epidemiology.py:667-834) -> t%02d_* arrays.  Per call: the look-back slices of contact
    rows exactly as the reference cuts them (start markers, off-by-one of :691-692 included), the traced ids, the keyed
    scan uniforms, and per primary contact -- in the reference's processing order -- the ordered list of draws it
    consumed (timesteps, uniforms, rounded log-normal days, selected secondary ids)."""
    N = rec.days[0]["n_agents_total"]
    out["trace_days"] = np.array([t["day"] for t in rec.traces], dtype=np.int32)
    for ti, t in enumerate(rec.traces):
        p = "t%02d_" % ti
        day, rows, markers = t["day"], t["rows"], t["markers"]
        traced = set(t["traced"])
        ev = t["events"]
        first_prim = next((i for i, e in enumerate(ev) if e[0] == "sq" and e[2] == 2), len(ev))
        start_ev = first_prim - 1 if first_prim < len(ev) else len(ev)      # the timestep draw before the first sq
        scan_us = [e[1] for e in ev[:start_ev] if e[0] == "u"]
        assert all(e[0] == "u" for e in ev[:start_ev])
        sl_off, sl_rows, keys, us = [0], [], [], []
        k = 0
        for back in range(days_back + 1):
            d = day - back
            lo, hi = 0, len(rows)
            if d in markers:
                lo = markers[d]
                if d + 1 in markers:
                    hi = markers[d + 1] - 1
                sl = rows[lo:hi]
            else:
                sl = rows[:0]
            for ri, (sct, a1, a2) in enumerate(sl):
                if (int(a1) in traced) != (int(a2) in traced):
                    keys.append((back << 32) | ri); us.append(scan_us[k]); k += 1
            sl_rows.append(sl); sl_off.append(sl_off[-1] + len(sl))
        assert k == len(scan_us), (k, len(scan_us))
        out[p + "scalars"] = np.array([day, int(t["testing_enabled"]), int(t["quarantine_enabled"]), t["n_locals"]], np.int64)
        out[p + "traced"] = np.array(sorted(traced), np.int32)
        out[p + "slice_off"] = np.array(sl_off, np.int64)
        out[p + "rows"] = np.concatenate(sl_rows).astype(np.int32).reshape(-1, 3)
        out[p + "scan_key"] = np.array(keys, np.uint64); out[p + "scan_u"] = np.array(us, np.float64)
        # per-primary draw records
        starts = [i - 1 for i, e in enumerate(ev) if e[0] == "sq" and e[2] == 2]
        prim, off, draws = [], [0], []
        for j, st in enumerate(starts):
            en = starts[j + 1] if j + 1 < len(starts) else len(ev)
            assert ev[st][0] == "ts"
            prim.append(ev[st + 1][1])
            for e in ev[st:en]:
                if e[0] in ("u", "ts", "ln"):
                    draws.append(float(e[1]))
                elif e[0] == "sel":
                    draws.extend(float(x) for x in e[1])
            off.append(len(draws))
        out[p + "prim_id"] = np.array(prim, np.int32); out[p + "prim_off"] = np.array(off, np.int32)
        out[p + "prim_draws"] = np.array(draws, np.float64)
        out[p + "secondary"] = np.array(sorted({e[1] for e in ev if e[0] == "sq" and e[2] == 3}), np.int32)
        out[p + "updated"] = np.array(sorted(t["updated"]), np.int32)
        assert sorted(t["updated"]) == sorted(set(prim) | {e[1] for e in ev if e[0] == "sq" and e[2] == 3})
        # static + records
        resc = np.full(N, -1, np.int32); present = np.zeros(N, np.uint8)
        for a, v in t["res_cellid"].items():
            if v is not None:
                resc[a] = v
        recs = {}
        for tag, epi in (("pre_", t["pre_epi"]), ("post_", t["post_epi"])):
            quar = np.full((N, 3), NONE, np.int32); test = np.full((N, 2), NONE, np.int32); tres = np.full((N, 2), NONE, np.int32)
            for a, e in epi.items():
                if tag == "pre_":
                    present[a] = 1
                for arr, key in ((quar, "quarantine_days"), (test, "test_day"), (tres, "test_result_day")):
                    v = e.get(key)
                    if v is not None and len(v) > 0:
                        arr[a, :len(v)] = v
            out[p + tag + "quar"], out[p + tag + "test"], out[p + tag + "tres"] = quar, test, tres
        out[p + "res_cellid"], out[p + "epi_present"] = resc, present
        # residences CSR: members = residents (minus nobody yet) then staff; accommodation from this call's snapshot
        cells = dict(rec.residences)
        for cid, mem in t["accom"].items():
            if cid in cells and cells[cid][1] == [] and cid not in getattr(rec, "_hh_inst", set()):
                cells[cid] = (mem, [])
        ncell = max(cells) + 1 if cells else 1
        roff = np.zeros(ncell + 1, np.int64); nres = np.zeros(ncell, np.int32); mem = []
        for c in range(ncell):
            r, sfx = cells.get(c, ([], []))
            mem.extend(r); mem.extend(sfx); nres[c] = len(r); roff[c + 1] = len(mem)
        out[p + "res_off"], out[p + "res_nres"], out[p + "res_members"] = roff, nres, np.array(mem, np.int32)


RUNS = [
    # name, epidemiology variant (None = data/epidemiology.json), seed, days, days kept in the fixture[, population]
    ("ref_1k_seed6", None, 6, 20, [1, 2, 5, 9, 12, 16, 20]),
    ("ref_1k_strict_seed7", "epidemiology_strict_days.json", 7, 25, [3, 10, 18, 21, 25]),
    # BASELINE.json configs 2 and 5 at their own size and parameter file: the 1k population tiled x10
    # (make_pop10k.py; the upstream 10k agents.json is missing), 90 days each, about 25 min per run on one core
    ("ref_10k_normal_seed11", "epidemiology_normal_days.json", 11, 90, [2, 15, 30, 45, 60, 75, 90], "10k"),
    ("ref_10k_strict_seed12", "epidemiology_strict_days.json", 12, 90, [3, 10, 20, 35, 50, 70, 90], "10k"),
]


def main():
    import subprocess
    if len(sys.argv) > 1 and sys.argv[1] == "--one":
        name, variant, seed, days = sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
        keep = [int(x) for x in sys.argv[6].split(",")]
        pop = sys.argv[7] if len(sys.argv) > 7 else "1k"
        variant = None if variant == "-" else variant
        if pop == "10k":
            import make_pop10k
            if not os.path.exists("/tmp/pop10k/agents.json"):
                make_pop10k.tile("/tmp/pop10k", 10)
            wd = rh.prepare_workdir("/tmp/refwork_" + name, popsubfolder="tiled10k", pop_src="/tmp/pop10k",
                                    epidemiology_variant=variant)
            # only the kept days are recorded (a 90-day record of every day would not fit the fixture budget)
            rec = rh.run_reference(wd, seed=seed, days=days, popsubfolder="tiled10k", record_days=set(keep))
        else:
            wd = rh.prepare_workdir("/tmp/refwork_" + name, epidemiology_variant=variant)
            rec = rh.run_reference(wd, seed=seed, days=days)
        out = encode_run(rec, keep)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        sys.stderr.write("%s: %d days recorded, %d kept, %d advance KATs\n" % (name, len(rec.days), len(keep),
                                                                               len(rec.advance_calls)))
        return
    only = sys.argv[2:] if len(sys.argv) > 2 and sys.argv[1] == "--runs" else None
    for run in RUNS:  # one interpreter per run: sim.py keeps module-level state
        name, variant, seed, days, keep = run[:5]
        if only is not None and name not in only:
            continue
        subprocess.check_call([sys.executable, os.path.abspath(__file__), "--one", name, variant or "-", str(seed),
                               str(days), ",".join(map(str, keep))] + list(run[5:]))


if __name__ == "__main__":
    main()
