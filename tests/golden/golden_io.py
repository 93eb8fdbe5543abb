"""Decoder for the golden fixtures written by make_golden.py (test infrastructure)."""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import mtd_oracle as mo  # noqa: E402

RUNS = {"ref_1k_seed6": "epidemiology.json", "ref_1k_strict_seed7": "epidemiology_strict_days.json",
        # hand-made cells through the reference's ContactNetwork (make_cells_kat.py): edge cases + one cell per K1 size class
        "ref_cells_kat": "epidemiology.json",
        # BASELINE.json configs 2 / 5 at their own size and parameter file: the 1k population tiled x10 (make_pop10k.py),
        # 90-day runs of the unmodified reference, seven kept days each
        "ref_10k_normal_seed11": "epidemiology_normal_days.json",
        "ref_10k_strict_seed12": "epidemiology_strict_days.json"}
NONE = -32768


class GoldenDay:
    pass


def _fill_epi(st: mo.AgentState, z, p, tag, ids):
    present = z[p + tag + "epi_present"]
    off = z[p + tag + "sch_off"]
    sch = z[p + tag + "sch"]
    for i, a in enumerate(ids):
        if not present[i]:
            continue
        ent = sch[off[i]:off[i + 1]]
        assert len(ent) <= mo.MAXSCHED
        st.sched_cnt[a] = len(ent)
        for j, (d, t, ts) in enumerate(ent):
            st.sched[a, j] = mo.pack_sched(d, t, ts)
    st.test_day[ids] = z[p + tag + "test"]
    st.test_result_day[ids] = z[p + tag + "tres"]
    st.quarantine[ids] = z[p + tag + "quar"]
    st.hosp[ids] = z[p + tag + "hosp"]


def load_run(name):
    z = np.load(os.path.join(HERE, name + ".npz"))
    return z


def load_day(z, day) -> GoldenDay:
    p = "d%02d_" % day
    g = GoldenDay()
    sc = z[p + "scalars"]
    g.day, g.weekday = int(sc[0]), int(sc[1])
    g.testing_enabled, g.quarantine_enabled, g.contact_tracing_enabled = bool(sc[2]), bool(sc[3]), bool(sc[4])
    g.N, g.n_locals = int(sc[5]), int(sc[6])
    g.mhd = float(z[p + "mhd"][0])
    g.iv_cell, g.iv_agent = z[p + "iv_cell"], z[p + "iv_agent"]
    g.iv_start, g.iv_end = z[p + "iv_start"], z[p + "iv_end"]
    g.cell_type, g.cell_indid = z["cell_type"], z["cell_indid"]
    g.contacts = z[p + "contacts"]
    g.exposed_ids = z[p + "exposed_ids"]
    ids = z[p + "act_ids"]
    g.act_ids = ids
    states = []
    for tag in ("pre_", "post_"):
        st = mo.AgentState(g.N)
        st.age_bracket[ids] = z[p + "age_bracket"]
        st.epi_age_bracket[ids] = z[p + "epi_age_bracket"]
        st.soc_rate[ids] = z[p + "soc_rate"]
        st.res_cellid[ids] = z[p + "res_cellid"]
        st.seir[:] = z[p + tag + "seir"]
        st.inf_type[:] = z[p + tag + "inf_type"]
        st.inf_sev[:] = z[p + tag + "inf_sev"]
        st.last_dose_day[:] = z[p + tag + "last_dose"]
        _fill_epi(st, z, p, tag, ids)
        states.append(st)
    g.pre, g.post = states
    ek = z[p + "expo_key"]
    # selected partner ids: the reference creates one contact row per pick, in pick order (contactnetwork.py:386-424),
    # and the tape's selection records are in call order -> walk both in step
    so, sa = z[p + "sel_off"], z[p + "sel_agent"]
    sel_ids = np.zeros(int(so[-1]) if len(so) else 0, np.int32)
    assert len(sel_ids) == len(g.contacts)
    for i in range(len(sa)):
        r = g.contacts[so[i]:so[i + 1]]
        assert np.all((r[:, 1] == sa[i]) | (r[:, 2] == sa[i]))
        sel_ids[so[i]:so[i + 1]] = np.where(r[:, 1] == sa[i], r[:, 2], r[:, 1])
    g.tape = mo.TapeArrays(z[p + "sel_cell"], z[p + "sel_agent"], z[p + "sel_off"], z[p + "sel_picks"],
                           z[p + "pair_key"], z[p + "pair_u"], ek[:, 3] if len(ek) else np.zeros(0, np.int32),
                           z[p + "expo_row"], sel_ids=sel_ids)
    g.expo_key = ek
    g.expo_ts_off, g.expo_ts = z[p + "expo_ts_off"], z[p + "expo_ts"]
    g.pair_key, g.pair_u = z[p + "pair_key"], z[p + "pair_u"]
    return g


def all_days():
    for name in RUNS:
        z = load_run(name)
        for d in z["days"]:
            yield name, int(d)


class GoldenTrace:
    pass


def load_trace(z, ti) -> GoldenTrace:
    """one recorded Epidemiology.contact_tracing call (epidemiology.py:667-834)"""
    p = "t%02d_" % ti
    t = GoldenTrace()
    sc = z[p + "scalars"]
    t.day, t.testing_enabled, t.quarantine_enabled, t.n_locals = int(sc[0]), bool(sc[1]), bool(sc[2]), int(sc[3])
    t.traced, t.slice_off, t.rows = z[p + "traced"], z[p + "slice_off"], z[p + "rows"]
    t.tape = mo.TraceTapeArrays(z[p + "scan_key"], z[p + "scan_u"], z[p + "prim_id"], z[p + "prim_off"], z[p + "prim_draws"])
    t.primary, t.secondary, t.updated = np.sort(z[p + "prim_id"]), z[p + "secondary"], z[p + "updated"]
    t.res_off, t.res_nres, t.res_members = z[p + "res_off"], z[p + "res_nres"], z[p + "res_members"]
    t.epi_present = z[p + "epi_present"]
    N = len(t.epi_present)
    states = []
    for tag in ("pre_", "post_"):
        st = mo.AgentState(N)
        st.res_cellid[:] = z[p + "res_cellid"]
        st.quarantine[:] = z[p + tag + "quar"]; st.test_day[:] = z[p + tag + "test"]
        st.test_result_day[:] = z[p + tag + "tres"]
        states.append(st)
    t.pre, t.post = states
    return t


def all_traces():
    for name in RUNS:
        z = load_run(name)
        if "trace_days" in z.files:
            for i in range(len(z["trace_days"])):
                yield name, i
