"""ctypes front-end of the CPU oracle (oracle/libmtd_oracle.so).  TEST INFRASTRUCTURE ONLY.

Used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs as the checker.
The product package (mtdcovabm_b200/) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
MAXSCHED = 12
NONE = -32768
EXPO_W = 21
MODE_REPLAY, MODE_PHILOX = 0, 1


def build(force=False):
    so = os.path.join(HERE, "libmtd_oracle.so")
    src = [os.path.join(HERE, f) for f in ("mtd_oracle.c", "mtd_oracle.h", "Makefile")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.check_call(["make", "-C", HERE, "-s", "-B", "libmtd_oracle.so"])
    return so


class Tape(C.Structure):
    _fields_ = [("n_sel", C.c_int64), ("sel_key", C.c_void_p), ("sel_off", C.c_void_p), ("sel_picks", C.c_void_p),
                ("n_pair", C.c_int64), ("pair_k1", C.c_void_p), ("pair_k2", C.c_void_p), ("pair_u", C.c_void_p),
                ("n_expo", C.c_int64), ("expo_agent", C.c_void_p), ("expo_row", C.c_void_p)]


class DayCtx(C.Structure):
    _fields_ = [("mode", C.c_int32), ("seed", C.c_uint32), ("day", C.c_int32), ("mhd", C.c_double),
                ("testing_enabled", C.c_int32), ("quarantine_enabled", C.c_int32), ("tape", C.POINTER(Tape))]


class Agents(C.Structure):
    _fields_ = [("n_agents", C.c_int64),
                ("age_bracket", C.c_void_p), ("epi_age_bracket", C.c_void_p), ("soc_rate", C.c_void_p),
                ("res_cellid", C.c_void_p),
                ("seir", C.c_void_p), ("inf_type", C.c_void_p), ("inf_sev", C.c_void_p), ("last_dose_day", C.c_void_p),
                ("sched", C.c_void_p), ("sched_cnt", C.c_void_p), ("test_day", C.c_void_p),
                ("test_result_day", C.c_void_p), ("quarantine", C.c_void_p), ("hosp", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [("pairs_evaluated", C.c_int64), ("n_cells", C.c_int64), ("n_potential", C.c_int64),
                ("tape_miss", C.c_int64), ("sched_overflow", C.c_int64)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.mtdo_contact_network.restype = C.c_int64
        _lib.mtdo_transmission.restype = C.c_int64
        _lib.mtdo_advance.restype = C.c_int64
        _lib.mtdo_np_pairwise_sum.restype = C.c_double
        _lib.mtdo_u53.restype = C.c_double
    return _lib


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


class AgentState:
    """Dense per-agent arrays (index = agent id); the same layout the C-ABI of the product uses."""

    STATIC = (("age_bracket", np.int8), ("epi_age_bracket", np.int8), ("soc_rate", np.float64), ("res_cellid", np.int32))
    DYN = (("seir", np.int8, ()), ("inf_type", np.int8, ()), ("inf_sev", np.int8, ()), ("last_dose_day", np.int32, ()),
           ("sched", np.int32, (MAXSCHED,)), ("sched_cnt", np.uint8, ()), ("test_day", np.int32, (2,)),
           ("test_result_day", np.int32, (2,)), ("quarantine", np.int32, (3,)), ("hosp", np.int32, (3,)))

    def __init__(self, n):
        self.n = int(n)
        self.age_bracket = np.zeros(n, np.int8)
        self.epi_age_bracket = np.zeros(n, np.int8)
        self.soc_rate = np.ones(n, np.float64)
        self.res_cellid = np.full(n, -1, np.int32)
        self.seir = np.zeros(n, np.int8)
        self.inf_type = np.zeros(n, np.int8)
        self.inf_sev = np.zeros(n, np.int8)
        self.last_dose_day = np.full(n, NONE, np.int32)
        self.sched = np.zeros((n, MAXSCHED), np.int32)
        self.sched_cnt = np.zeros(n, np.uint8)
        self.test_day = np.full((n, 2), NONE, np.int32)
        self.test_result_day = np.full((n, 2), NONE, np.int32)
        self.quarantine = np.full((n, 3), NONE, np.int32)
        self.hosp = np.full((n, 3), NONE, np.int32)

    def copy(self):
        o = AgentState.__new__(AgentState)
        o.n = self.n
        for name, *_ in self.STATIC + self.DYN:
            setattr(o, name, getattr(self, name).copy())
        return o

    def struct(self):
        s = Agents()
        s.n_agents = self.n
        for name, *_ in self.STATIC + self.DYN:
            a = getattr(self, name)
            assert a.flags["C_CONTIGUOUS"]
            setattr(s, name, a.ctypes.data)
        return s

    def dyn_equal(self, other):
        return {name: bool(np.array_equal(getattr(self, name), getattr(other, name))) for name, *_ in self.DYN}


def pack_sched(day, trans, ts):
    return np.int32(np.uint32((int(day) & 0xFFFF) << 16 | (int(trans) & 0xFF) << 8 | (int(ts) & 0xFF)))


def unpack_sched(v):
    v = int(np.int32(v))
    return v >> 16, (v >> 8) & 0xFF, v & 0xFF


class TapeArrays:
    """Sorted key/value arrays of one day's replay tape."""

    def __init__(self, sel_cell, sel_agent, sel_off, sel_picks, pair_key, pair_u, expo_agent, expo_row, sel_ids=None):
        sk = (sel_cell.astype(np.uint64) << np.uint64(32)) | sel_agent.astype(np.uint32).astype(np.uint64)
        order = np.argsort(sk, kind="stable")
        self.sel_key = np.ascontiguousarray(sk[order])
        lens = np.diff(sel_off)[order]
        self.sel_off = np.zeros(len(sk) + 1, np.int32)
        self.sel_off[1:] = np.cumsum(lens)
        self.sel_picks = np.concatenate([sel_picks[sel_off[i]:sel_off[i + 1]] for i in order]).astype(np.int32) \
            if len(order) else np.zeros(0, np.int32)
        # the same selections as partner AGENT IDS (what the product's tape holds; the oracle keeps the indices into the
        # reference-ordered candidate list, which also pins its restatement of the dict insertion order)
        self.sel_ids = None
        if sel_ids is not None:
            self.sel_ids = np.concatenate([sel_ids[sel_off[i]:sel_off[i + 1]] for i in order]).astype(np.int32) \
                if len(order) else np.zeros(0, np.int32)
        k1 = (pair_key[:, 0].astype(np.uint64) << np.uint64(32)) | pair_key[:, 1].astype(np.uint32).astype(np.uint64)
        k2 = pair_key[:, 2].astype(np.uint32)
        po = np.lexsort((k2, k1))
        self.pair_k1 = np.ascontiguousarray(k1[po])
        self.pair_k2 = np.ascontiguousarray(k2[po])
        self.pair_u = np.ascontiguousarray(pair_u[po].astype(np.float64))
        eo = np.argsort(expo_agent, kind="stable")
        self.expo_agent = np.ascontiguousarray(expo_agent[eo].astype(np.int32))
        self.expo_row = np.ascontiguousarray(expo_row[eo].astype(np.float64)).reshape(-1, EXPO_W)

    def struct(self):
        t = Tape()
        t.n_sel = len(self.sel_key); t.sel_key = self.sel_key.ctypes.data
        t.sel_off = self.sel_off.ctypes.data; t.sel_picks = self.sel_picks.ctypes.data
        t.n_pair = len(self.pair_k1); t.pair_k1 = self.pair_k1.ctypes.data
        t.pair_k2 = self.pair_k2.ctypes.data; t.pair_u = self.pair_u.ctypes.data
        t.n_expo = len(self.expo_agent); t.expo_agent = self.expo_agent.ctypes.data
        t.expo_row = self.expo_row.ctypes.data
        return t


def make_ctx(mode, day, mhd, testing_enabled, quarantine_enabled, seed=0, tape: TapeArrays | None = None):
    ctx = DayCtx()
    ctx.mode = mode; ctx.seed = seed; ctx.day = day; ctx.mhd = mhd
    ctx.testing_enabled = int(testing_enabled); ctx.quarantine_enabled = int(quarantine_enabled)
    keep = None
    if tape is not None:
        keep = tape.struct()
        ctx.tape = C.pointer(keep)
    ctx._keep = (keep, tape)
    return ctx


def contact_network(params, ctx, iv_cell, iv_agent, iv_start, iv_end, cell_type, cell_indid, agents: AgentState,
                    cap=None):
    """-> (rows[D,5] int32, row_dur[D], row_cell[D], Stats)"""
    M = len(iv_cell)
    cap = int(cap if cap is not None else max(1024, 12 * M))
    rows = np.zeros((cap, 5), np.int32); dur = np.zeros(cap, np.int32); rc = np.zeros(cap, np.int32)
    st = Stats()
    a = agents.struct()
    iv_cell = np.ascontiguousarray(iv_cell, np.int32); iv_agent = np.ascontiguousarray(iv_agent, np.int32)
    iv_start = np.ascontiguousarray(iv_start, np.uint8); iv_end = np.ascontiguousarray(iv_end, np.uint8)
    cell_type = np.ascontiguousarray(cell_type, np.int8); cell_indid = np.ascontiguousarray(cell_indid, np.int16)
    n = lib().mtdo_contact_network(C.byref(params), C.byref(ctx), C.c_int64(M), _p(iv_cell), _p(iv_agent), _p(iv_start),
                                   _p(iv_end), _p(cell_type), _p(cell_indid), C.byref(a), C.c_int64(cap), _p(rows),
                                   _p(dur), _p(rc), C.byref(st))
    if n < 0:
        raise RuntimeError("oracle: contact row capacity exceeded")
    return rows[:n].copy(), dur[:n].copy(), rc[:n].copy(), st


def transmission(params, ctx, iv_cell, iv_agent, iv_start, iv_end, cell_type, rows, row_dur, row_cell,
                 agents: AgentState, want_p=False):
    """Mutates `agents`.  -> (exposed ids in order, p per row or None, Stats)"""
    M = len(iv_cell); n_rows = len(rows)
    exposed = np.zeros(max(16, n_rows), np.int32)
    p = np.zeros(max(1, n_rows), np.float64) if want_p else None
    st = Stats()
    a = agents.struct()
    iv_cell = np.ascontiguousarray(iv_cell, np.int32); iv_agent = np.ascontiguousarray(iv_agent, np.int32)
    iv_start = np.ascontiguousarray(iv_start, np.uint8); iv_end = np.ascontiguousarray(iv_end, np.uint8)
    cell_type = np.ascontiguousarray(cell_type, np.int8)
    rows = np.ascontiguousarray(rows, np.int32); row_dur = np.ascontiguousarray(row_dur, np.int32)
    row_cell = np.ascontiguousarray(row_cell, np.int32)
    n = lib().mtdo_transmission(C.byref(params), C.byref(ctx), C.c_int64(M), _p(iv_cell), _p(iv_agent), _p(iv_start),
                                _p(iv_end), _p(cell_type), C.c_int64(n_rows), _p(rows), _p(row_dur), _p(row_cell),
                                C.byref(a), C.c_int64(len(exposed)), _p(exposed), _p(p), C.byref(st))
    if n < 0:
        raise RuntimeError("oracle: exposed capacity exceeded")
    return exposed[:n].copy(), (p[:n_rows] if want_p else None), st


def advance(day, agents: AgentState, mask=None):
    a = agents.struct()
    if mask is not None:
        mask = np.ascontiguousarray(mask, np.uint8)
    return lib().mtdo_advance(C.c_int32(day), C.byref(a), _p(mask))


def build_intervals(ent_off, ent_ts, ent_cell, grp_off, grp_agent, grp_rescell, cap):
    """itinerary.py:1403-1488 -> (iv_agent, iv_cell, iv_start, iv_end) in emission order"""
    a, c = np.zeros(cap, np.int32), np.zeros(cap, np.int32)
    s, e = np.zeros(cap, np.uint8), np.zeros(cap, np.uint8)
    L = lib()
    L.mtdo_build_intervals.restype = C.c_int64
    n = L.mtdo_build_intervals(C.c_int64(len(ent_off) - 1), _p(ent_off), _p(ent_ts), _p(ent_cell), _p(grp_off), _p(grp_agent),
                               _p(grp_rescell), C.c_int64(cap), _p(a), _p(c), _p(s), _p(e))
    if n < 0:
        raise RuntimeError("oracle: interval capacity exceeded")
    return a[:n].copy(), c[:n].copy(), s[:n].copy(), e[:n].copy()


def schedule_vaccinations(ctx, n_locals, k, last_dose_day, vacc_day):
    """epidemiology.py:911-976 (device part): mutates vacc_day [n][2] -> ascending ids scheduled"""
    sel = np.zeros(len(last_dose_day), np.uint8)
    L = lib()
    L.mtdo_schedule_vaccinations.restype = C.c_int64
    L.mtdo_schedule_vaccinations(C.byref(ctx), C.c_int64(n_locals), C.c_int64(k), _p(last_dose_day), _p(vacc_day), _p(sel))
    return np.nonzero(sel)[0].astype(np.int32)


def vaccination_day(day, n_locals, last_dose_day, vacc_day):
    L = lib()
    L.mtdo_vaccination_day.restype = None
    L.mtdo_vaccination_day(C.c_int32(day), C.c_int64(n_locals), _p(last_dose_day), _p(vacc_day))


def set_log_ulp_shift(k):
    """test knob: move the oracle's two per-cell logarithms by k ulps (0 = glibc's value)"""
    C.c_int.in_dll(lib(), "mtdo_log_ulp_shift").value = int(k)


def glibc_log(x):
    x = np.ascontiguousarray(x, np.float64)
    out = np.empty_like(x)
    lib().mtdo_log_array(C.c_int64(len(x)), _p(x), _p(out))
    return out


def process_test_results(params, ctx, agents: AgentState, epi_present=None, replay_agents=None, replay_u=None):
    """itinerary.py:2283-2528 (record bookkeeping) over all agents -> ids added to contact_tracing_agent_ids (ascending).
    Replay mode: replay_agents ascending ids, replay_u [n, 2] the result draw and the quarantine draw of each."""
    a = agents.struct()
    pos = np.zeros(agents.n, np.uint8)
    if epi_present is not None:
        epi_present = np.ascontiguousarray(epi_present, np.uint8)
    ra = None if replay_agents is None else np.ascontiguousarray(replay_agents, np.int32)
    ru = None if replay_u is None else np.ascontiguousarray(replay_u, np.float64)
    miss = C.c_int64(0)
    L = lib()
    L.mtdo_process_interventions.restype = C.c_int64
    L.mtdo_process_interventions(C.byref(params), C.byref(ctx), C.byref(a), _p(epi_present), C.c_int64(0 if ra is None else len(ra)),
                                 _p(ra), _p(ru), _p(pos), C.byref(miss))
    process_test_results.tape_miss = miss.value
    return np.nonzero(pos)[0].astype(np.int32)


def seed_exposures(params, ctx, agents: AgentState, agent_ids, exposure_day=None, pre_exposure_days=0, timesteps=None,
                   asym_mult=None, draws=None):
    """itinerary.py:373-395 -> epidemiology.py:233-500 with reset_states: mutates `agents`; -> (out [n, 8], Stats)"""
    a = agents.struct()
    ids = np.ascontiguousarray(agent_ids, np.int32)
    ts = np.ascontiguousarray(np.arange(144) if timesteps is None else timesteps, np.uint8)
    ed = None if exposure_day is None else np.ascontiguousarray(exposure_day, np.int32)
    am = None if asym_mult is None else np.ascontiguousarray(asym_mult, np.float64)
    dr = None if draws is None else np.ascontiguousarray(draws, np.float64)
    out = np.zeros((len(ids), 8), np.int32)
    st = Stats()
    L = lib()
    L.mtdo_seed_exposures.restype = C.c_int64
    r = L.mtdo_seed_exposures(C.byref(params), C.byref(ctx), C.byref(a), C.c_int64(len(ids)), _p(ids), _p(ed),
                              C.c_int32(pre_exposure_days), _p(ts), C.c_int32(len(ts)), _p(am), _p(dr), _p(out), C.byref(st))
    if r != 0:
        raise ValueError("seed_exposures: agent id out of range")
    return out, st


def mask_intervals(day, iv_agent, iv_cell, agents: AgentState):
    """keep flags of the day's intervals under the intervention mask (deceased / quarantined / hospitalised agents stay
    in their residence cell only)"""
    a = agents.struct()
    keep = np.zeros(len(iv_agent), np.uint8)
    L = lib()
    L.mtdo_mask_intervals.restype = C.c_int64
    L.mtdo_mask_intervals(C.c_int32(day), C.c_int64(len(iv_agent)), _p(iv_agent), _p(iv_cell), C.byref(a), _p(keep))
    return keep.astype(bool)


# ---------------------------------------------------------------------------------- contact tracing
class TraceTape(C.Structure):
    _fields_ = [("n_scan", C.c_int64), ("scan_key", C.c_void_p), ("scan_u", C.c_void_p),
                ("n_prim", C.c_int64), ("prim_id", C.c_void_p), ("prim_off", C.c_void_p), ("prim_draws", C.c_void_p)]


class TraceStats(C.Structure):
    _fields_ = [("n_primary", C.c_int64), ("n_secondary", C.c_int64), ("n_scan_draws", C.c_int64),
                ("tape_miss", C.c_int64)]


class TraceTapeArrays:
    """replay tape of one contact_tracing call (tests/golden/make_golden.py:encode_traces)"""

    def __init__(self, scan_key, scan_u, prim_id, prim_off, prim_draws):
        o = np.argsort(scan_key, kind="stable")
        self.scan_key = np.ascontiguousarray(np.asarray(scan_key, np.uint64)[o])
        self.scan_u = np.ascontiguousarray(np.asarray(scan_u, np.float64)[o])
        self.prim_id = np.ascontiguousarray(prim_id, np.int32)
        self.prim_off = np.ascontiguousarray(prim_off, np.int32)
        self.prim_draws = np.ascontiguousarray(prim_draws, np.float64)

    def struct(self):
        t = TraceTape()
        t.n_scan = len(self.scan_key); t.scan_key = self.scan_key.ctypes.data; t.scan_u = self.scan_u.ctypes.data
        t.n_prim = len(self.prim_id); t.prim_id = self.prim_id.ctypes.data; t.prim_off = self.prim_off.ctypes.data
        t.prim_draws = self.prim_draws.ctypes.data
        return t


def contact_tracing(params, ctx, tape, traced, slice_off, rows3, res_off, res_nres, res_members, agents: AgentState,
                    epi_present, n_locals):
    """Mutates agents.quarantine / test_day / test_result_day.  -> (primary flags, secondary flags, TraceStats)"""
    L = lib()
    L.mtdo_contact_tracing.restype = C.c_int64
    N = agents.n
    prim = np.zeros(N, np.uint8); sec = np.zeros(N, np.uint8)
    st = TraceStats()
    a = agents.struct()
    traced = np.ascontiguousarray(traced, np.int32); slice_off = np.ascontiguousarray(slice_off, np.int64)
    rows3 = np.ascontiguousarray(rows3, np.int32)
    res_off = np.ascontiguousarray(res_off, np.int64); res_nres = np.ascontiguousarray(res_nres, np.int32)
    res_members = np.ascontiguousarray(res_members, np.int32)
    epi_present = np.ascontiguousarray(epi_present, np.uint8)
    ts = tape.struct() if tape is not None else TraceTape()
    L.mtdo_contact_tracing(C.byref(params), C.byref(ctx), C.byref(ts), C.c_int64(len(traced)), _p(traced),
                           C.c_int64(len(slice_off) - 1), _p(slice_off), _p(rows3), C.c_int64(len(res_nres)),
                           _p(res_off), _p(res_nres), _p(res_members), C.byref(a), _p(epi_present), C.c_int32(n_locals),
                           _p(prim), _p(sec), C.byref(st))
    return prim, sec, st
