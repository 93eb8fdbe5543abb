"""Reference containers <-> dense arrays.

The reference keeps per-agent state in Python dict-likes keyed by agent id (`vars_util.agents_seir_state`,
`agents_epi[id][...]`, vars.py:44-56, agentsutil.py:56-62) and static attributes behind `agents_static.get(id, name)`
(static.py:351-363).  The device works on dense arrays indexed by agent id.  This module flattens one into the other at
the boundary and scatters the results back; it holds no model logic.
"""
from __future__ import annotations

import numpy as np

from . import capi
from .epidemiologyclasses import InfectionType, SEIRState, Severity
from .params import MAXSCHED, NONE

_handles = {}


def get_handle(n_agents, n_cells, max_intervals, params, device=0):
    """one persistent handle per (population size, device): the reference builds a new ContactNetwork every day
    (contactnetwork_mp.py:287), the device buffers are kept across days"""
    key = (int(n_agents), int(n_cells), int(device))
    h = _handles.get(key)
    if h is None or h.max_intervals < max_intervals:
        if h is not None:
            h.close()
        h = capi.Handle(n_agents, n_cells, max(int(max_intervals * 1.5), 4096), device=device)
        _handles[key] = h
    h.set_params(params)
    return h


def close_handles():
    for h in _handles.values():
        h.close()
    _handles.clear()


def pack_sched(day, trans, ts):
    if not -32768 <= int(day) <= 32767:      # 16-bit day field of the packed schedule entry (MTD_NONE = -32768 is "None")
        raise ValueError("schedule day %d does not fit the 16-bit field of the packed entry" % int(day))
    v = (int(day) & 0xFFFF) << 16 | (int(trans) & 0xFF) << 8 | (int(ts) & 0xFF)
    return np.int32(v - (1 << 32) if v >= (1 << 31) else v)


def unpack_sched(v):
    v = int(np.int32(v))
    return [v >> 16, (v >> 8) & 0xFF, v & 0xFF]


class DenseState:
    """dense per-agent arrays in the layout of the C ABI (mtd_set_state / mtd_get_state)"""

    def __init__(self, n):
        self.n = n
        self.age_bracket = np.zeros(n, np.int8); self.epi_age_bracket = np.zeros(n, np.int8)
        self.soc_rate = np.ones(n, np.float64); self.res_cellid = np.full(n, -1, np.int32)
        self.seir = np.zeros(n, np.int8); self.inf_type = np.zeros(n, np.int8); self.inf_sev = np.zeros(n, np.int8)
        self.last_dose_day = np.full(n, NONE, np.int32)
        self.sched = np.zeros((n, MAXSCHED), np.int32); self.sched_cnt = np.zeros(n, np.uint8)
        self.test_day = np.full((n, 2), NONE, np.int32); self.test_result_day = np.full((n, 2), NONE, np.int32)
        self.quarantine = np.full((n, 3), NONE, np.int32); self.hosp = np.full((n, 3), NONE, np.int32)


def _has(v):
    return v is not None and len(v) > 0


def _rec(arr, i, v):
    if v is not None and len(v) > 0:
        arr[i, :len(v)] = [int(x) for x in v]


def flatten_intervals(cells_agents_timesteps):
    """{cellid: [(agent, start, end), ...]} in dict order = emission order (itinerary.py:1446-1452) ->
    iv_cell i32[M], iv_agent i32[M], iv_start u8[M], iv_end u8[M]"""
    import itertools
    n_c = len(cells_agents_timesteps)
    lens = np.fromiter((len(l) for l in cells_agents_timesteps.values()), np.int64, n_c)
    M = int(lens.sum())
    keys = np.fromiter(cells_agents_timesteps.keys(), np.int64, n_c)
    flat = np.fromiter(itertools.chain.from_iterable(itertools.chain.from_iterable(cells_agents_timesteps.values())),
                       np.int64, 3 * M).reshape(M, 3)
    return (np.repeat(keys, lens).astype(np.int32), flat[:, 0].astype(np.int32), flat[:, 1].astype(np.uint8),
            flat[:, 2].astype(np.uint8))


# Static attributes of the LOCAL agents never change during a run (static.py:351-363; only tourists are (re)written, on
# arrival, tourism.py:193-200,302-309), and the reference keeps one Static object for the whole run while it builds a new
# ContactNetwork every day: their dense copy is kept per Static object and filled as agents become active.
_static_cache = {}


def clear_static_cache():
    _static_cache.clear()


def _static_arrays(n_agents, agents_static, active_ids, n_locals):
    ent = _static_cache.get(id(agents_static))
    if ent is None or ent[0] is not agents_static or len(ent[1]) != n_agents:
        if len(_static_cache) >= 4:
            _static_cache.clear()
        ent = (agents_static, np.zeros(n_agents, bool), np.zeros(n_agents, np.int8), np.zeros(n_agents, np.int8),
               np.ones(n_agents, np.float64), np.full(n_agents, -1, np.int32))
        _static_cache[id(agents_static)] = ent
    _, have, age, epi, soc, res = ent
    active_ids = np.asarray(active_ids, np.int64)
    todo = active_ids[~have[active_ids]]
    get = agents_static.get
    for a in todo.tolist():
        age[a] = int(get(a, "age_bracket_index")); epi[a] = int(get(a, "epi_age_bracket_index"))
        soc[a] = float(get(a, "soc_rate")); res[a] = int(get(a, "res_cellid"))
    if n_locals is not None:
        have[todo[todo < int(n_locals)]] = True
    return age, epi, soc, res


def _dict_to(arr, d, conv=int):
    n = len(d)
    if n:
        arr[np.fromiter(d.keys(), np.int64, n)] = np.fromiter(map(conv, d.values()), np.int64, n)


def flatten(n_agents, agents_static, agents_epi, vars_util, active_ids, n_locals=None):
    """-> DenseState.  Dynamic state for every agent that has a SEIR state; static attributes for `active_ids` (the
    agents that appear in today's cells: the only ones whose attributes the path reads).  n_locals: ids below it are
    local agents, whose static attributes are cached across calls (None = no caching)."""
    st = DenseState(n_agents)
    st.age_bracket, st.epi_age_bracket, st.soc_rate, st.res_cellid = _static_arrays(n_agents, agents_static, active_ids, n_locals)
    _dict_to(st.seir, vars_util.agents_seir_state)
    _dict_to(st.inf_type, vars_util.agents_infection_type)
    _dict_to(st.inf_sev, vars_util.agents_infection_severity)
    for a, doses in getattr(vars_util, "agents_vaccination_doses", {}).items():
        if doses is not None and len(doses) > 0:
            st.last_dose_day[int(a)] = int(doses[-1])           # util.py:126: only the last dose is used
    epi_get = agents_epi.get if hasattr(agents_epi, "get") else agents_epi.__getitem__
    sched, sched_cnt = st.sched, st.sched_cnt
    for a in np.asarray(active_ids).tolist():          # plain Python from here on: one record of five optional lists per agent
        e = epi_get(a)
        if e is None:
            continue
        get = e.get
        ent = get("state_transition_by_day")
        if ent is not None and len(ent):
            if len(ent) > MAXSCHED:
                raise RuntimeError("agent %d holds %d schedule entries (> MTD_MAXSCHED)" % (a, len(ent)))
            sched_cnt[a] = len(ent)
            for j, (d, t, ts) in enumerate(ent):
                sched[a, j] = pack_sched(d, t, ts)
        v = get("test_day")
        if v is not None and len(v):
            _rec(st.test_day, a, v)
        v = get("test_result_day")
        if v is not None and len(v):
            _rec(st.test_result_day, a, v)
        v = get("quarantine_days")
        if v is not None and len(v):
            _rec(st.quarantine, a, v)
        v = get("hospitalisation_days")
        if v is not None and len(v):
            _rec(st.hosp, a, v)
    return st


def flatten_reference_loop(n_agents, agents_static, agents_epi, vars_util, active_ids):
    """the plain per-element restatement of flatten() (kept as the checker of the vectorised one, tests/test_bridge.py)"""
    st = DenseState(n_agents)
    for a in active_ids:
        a = int(a)
        st.age_bracket[a] = int(agents_static.get(a, "age_bracket_index"))
        st.epi_age_bracket[a] = int(agents_static.get(a, "epi_age_bracket_index"))
        st.soc_rate[a] = float(agents_static.get(a, "soc_rate"))
        st.res_cellid[a] = int(agents_static.get(a, "res_cellid"))
    for a, s in vars_util.agents_seir_state.items():
        st.seir[int(a)] = int(s)
    for a, v in vars_util.agents_infection_type.items():
        st.inf_type[int(a)] = int(v)
    for a, v in vars_util.agents_infection_severity.items():
        st.inf_sev[int(a)] = int(v)
    for a, doses in getattr(vars_util, "agents_vaccination_doses", {}).items():
        if doses is not None and len(doses) > 0:
            st.last_dose_day[int(a)] = int(doses[-1])
    for a in active_ids:
        a = int(a)
        e = agents_epi.get(a) if hasattr(agents_epi, "get") else agents_epi[a]
        if e is None:
            continue
        ent = e.get("state_transition_by_day") or []
        st.sched_cnt[a] = len(ent)
        for j, (d, t, ts) in enumerate(ent):
            st.sched[a, j] = pack_sched(d, t, ts)
        _rec(st.test_day, a, e.get("test_day")); _rec(st.test_result_day, a, e.get("test_result_day"))
        _rec(st.quarantine, a, e.get("quarantine_days")); _rec(st.hosp, a, e.get("hospitalisation_days"))
    return st


def upload(h, st: DenseState, cell_type, cell_indid):
    h.set_cells(cell_type, cell_indid)
    h.set_static(st.age_bracket, st.epi_age_bracket, st.soc_rate, st.res_cellid)
    h.set_state(seir=st.seir, inf_type=st.inf_type, inf_sev=st.inf_sev, last_dose_day=st.last_dose_day, sched=st.sched,
                sched_cnt=st.sched_cnt, test_day=st.test_day, test_result_day=st.test_result_day,
                quarantine=st.quarantine, hosp=st.hosp)


def _lst(row):
    return None if row[0] == NONE else [int(x) for x in row]


def scatter(dev: dict, ids, agents_epi, vars_util):
    """write the device state of `ids` back into the reference's containers (epidemiology.py:215-225)"""
    for a in ids:
        a = int(a)
        e = agents_epi[a]
        cnt = int(dev["sched_cnt"][a])
        e["state_transition_by_day"] = [unpack_sched(v) for v in dev["sched"][a, :cnt]]
        e["test_day"] = _lst(dev["test_day"][a]); e["test_result_day"] = _lst(dev["test_result_day"][a])
        e["quarantine_days"] = _lst(dev["quarantine"][a]); e["hospitalisation_days"] = _lst(dev["hosp"][a])
        vars_util.agents_seir_state[a] = SEIRState(int(dev["seir"][a]))
        vars_util.agents_infection_type[a] = InfectionType(int(dev["inf_type"][a]))
        vars_util.agents_infection_severity[a] = Severity(int(dev["inf_sev"][a]))
