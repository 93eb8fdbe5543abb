"""Shared helpers of the parity tests (test infrastructure; may import oracle/)."""
from __future__ import annotations

import numpy as np

from mtdcovabm_b200 import capi
from mtdcovabm_b200 import params as prm
from oracle import mtd_oracle as mo

STATE_FIELDS = ("seir", "inf_type", "inf_sev", "sched", "sched_cnt", "test_day", "test_result_day", "quarantine", "hosp")


def make_handle(P, st: mo.AgentState, cell_type, cell_indid, max_intervals, **kw):
    h = capi.Handle(st.n, len(cell_type), max_intervals, **kw)
    h.set_params(P)
    h.set_cells(np.ascontiguousarray(cell_type, np.int8), np.ascontiguousarray(cell_indid, np.int16))
    h.set_static(st.age_bracket, st.epi_age_bracket, st.soc_rate, st.res_cellid)
    upload_state(h, st)
    return h


def upload_state(h, st: mo.AgentState):
    h.set_state(seir=st.seir, inf_type=st.inf_type, inf_sev=st.inf_sev, last_dose_day=st.last_dose_day, sched=st.sched,
                sched_cnt=st.sched_cnt, test_day=st.test_day, test_result_day=st.test_result_day,
                quarantine=st.quarantine, hosp=st.hosp)


def state_diff(dev: dict, st: mo.AgentState):
    """names of the per-agent arrays that differ between the device state and an oracle AgentState"""
    bad = []
    for f in STATE_FIELDS:
        a, b = dev[f], getattr(st, f)
        if f == "sched":      # only the first sched_cnt entries are meaningful
            cnt = st.sched_cnt.astype(np.int64)
            mask = np.arange(a.shape[1])[None, :] < cnt[:, None]
            if not np.array_equal(a[mask], b[mask]):
                bad.append(f)
        elif not np.array_equal(a, b):
            bad.append(f)
    return bad


def set_tape(h, tape: mo.TapeArrays):
    h.set_tape(tape.sel_key, tape.sel_off, tape.sel_ids, tape.pair_k1, tape.pair_k2, tape.pair_u, tape.expo_agent,
               tape.expo_row)


def tile_day(g, reps, agent_major=True, seed=0, infectious_frac=0.02):
    """Replicate a golden day `reps` times (agent id += r*N, cell id += r*C) -> synthetic population
    (SURVEY.md section 8d).  Returns dict with stream arrays, cell tables and an oracle AgentState."""
    N1, C1 = g.N, len(g.cell_type)
    M1 = len(g.iv_cell)
    r = np.repeat(np.arange(reps, dtype=np.int64), M1)
    iv_agent = (np.tile(g.iv_agent.astype(np.int64), reps) + r * N1).astype(np.int32)
    iv_cell = (np.tile(g.iv_cell.astype(np.int64), reps) + r * C1).astype(np.int32)
    iv_start = np.tile(g.iv_start, reps); iv_end = np.tile(g.iv_end, reps)
    if agent_major:   # emission order of an itinerary stage that walks agents (itinerary.py:1403-1452)
        o = np.argsort(iv_agent, kind="stable")
        iv_agent, iv_cell, iv_start, iv_end = iv_agent[o], iv_cell[o], iv_start[o], iv_end[o]
    st = mo.AgentState(N1 * reps)
    for name in ("age_bracket", "epi_age_bracket", "soc_rate"):
        getattr(st, name)[:] = np.tile(getattr(g.pre, name), reps)
    rc = np.tile(g.pre.res_cellid.astype(np.int64), reps) + np.repeat(np.arange(reps, dtype=np.int64), N1) * C1
    rc[np.tile(g.pre.res_cellid, reps) < 0] = -1
    st.res_cellid[:] = rc.astype(np.int32)
    rng = np.random.default_rng(seed)
    present = np.tile(g.pre.seir != 0, reps)
    st.seir[:] = np.where(present, 1, 0)
    inf = present & (rng.random(N1 * reps) < infectious_frac)
    st.seir[inf] = 3; st.inf_type[inf] = 3; st.inf_sev[inf] = 1
    return dict(iv_agent=np.ascontiguousarray(iv_agent), iv_cell=np.ascontiguousarray(iv_cell),
                iv_start=np.ascontiguousarray(iv_start), iv_end=np.ascontiguousarray(iv_end),
                cell_type=np.tile(g.cell_type, reps), cell_indid=np.tile(g.cell_indid, reps), state=st)
