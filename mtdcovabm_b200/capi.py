"""ctypes binding of libmtd_b200.so (C ABI declared in include/mtd_b200.h).

There is no CPU fallback: if the CUDA library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .params import EXPO_W, MAXSCHED, NONE, ParamsStruct

PKG = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(PKG, "libmtd_b200.so")
MEM_HOST, MEM_DEVICE, MEM_STAGED, MEM_BUILT = 0, 1, 2, 3
MODE_REPLAY, MODE_PHILOX = 0, 1
MAX_CELL_INTERVALS = 16384     # include/mtd_b200.h MTD_MAX_CELL_INTERVALS
NOALLOC_MAX_AGENTS = 1 << 26   # include/mtd_b200.h MTD_NOALLOC_MAX_AGENTS
DELTA_WORDS = 20
APPLY_NONE, APPLY_LOCAL, APPLY_EXCHANGE = 0, 1, 2
IPC_HANDLE_BYTES = 64

EXPORTS = ["mtd_create", "mtd_destroy", "mtd_last_error", "mtd_abi_version", "mtd_set_params", "mtd_set_cells",
           "mtd_set_static", "mtd_set_state", "mtd_get_state", "mtd_set_tape", "mtd_simulate_contact_network",
           "mtd_simulate_direct_contacts", "mtd_apply_deltas", "mtd_advance_seir", "mtd_get_counters",
           "mtd_get_contacts", "mtd_get_exposed", "mtd_get_infection_prob", "mtd_occupancy", "mtd_sync",
           "mtd_set_profiling", "mtd_get_stage_ms", "mtd_set_export", "mtd_prefetch_intervals",
           "mtd_export_day_async", "mtd_export_wait", "mtd_set_residences", "mtd_contact_tracing", "mtd_get_traced",
           "mtd_exchange_init", "mtd_exchange_open", "mtd_exchange_ready", "mtd_set_contacts", "mtd_seed_exposures", "mtd_set_intervention_mask", "mtd_set_k0_overlap",
           "mtd_process_test_results", "mtd_debug_log", "mtd_set_k6_tape", "mtd_contact_tracing_scan", "mtd_contact_tracing_finish", "mtd_enable_vaccination", "mtd_schedule_vaccinations", "mtd_set_vaccination",
           "mtd_get_vaccination", "mtd_build_intervals", "mtd_get_intervals", "mtd_occupancy_noalloc"]
STAGES = ("h2d", "k0_csr", "k1_contacts", "k2_rows", "k3_eval", "k3_apply", "k4_advance")


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("n_agents", C.c_int64), ("n_cells", C.c_int64), ("max_intervals", C.c_int64),
                ("max_rows", C.c_int64), ("max_exposures", C.c_int64), ("stream", C.c_void_p)]


class Day(C.Structure):
    _fields_ = [("mode", C.c_int32), ("seed", C.c_uint32), ("day", C.c_int32), ("mhd", C.c_double),
                ("testing_enabled", C.c_int32), ("quarantine_enabled", C.c_int32)]


class Tape(C.Structure):
    _fields_ = [("n_sel", C.c_int64), ("sel_key", C.c_void_p), ("sel_off", C.c_void_p), ("sel_picks", C.c_void_p),
                ("n_pair", C.c_int64), ("pair_k1", C.c_void_p), ("pair_k2", C.c_void_p), ("pair_u", C.c_void_p),
                ("n_expo", C.c_int64), ("expo_agent", C.c_void_p), ("expo_row", C.c_void_p)]


class Counters(C.Structure):
    _fields_ = [("n_slots", C.c_int64), ("n_rows", C.c_int64), ("n_exposed", C.c_int64), ("pairs_evaluated", C.c_int64),
                ("n_potential", C.c_int64), ("n_si_rows", C.c_int64), ("sched_overflow", C.c_int64),
                ("tape_miss", C.c_int64), ("class_cells", C.c_int64 * 9), ("n_incidences", C.c_int64), ("error", C.c_int64)]


class DayResult(C.Structure):
    """mtd_day_result (include/mtd_b200.h): counts + occupancy of one exported day"""
    _fields_ = [("day", C.c_int64), ("n_rows", C.c_int64), ("n_exposed", C.c_int64), ("occupancy", C.c_int64 * 6),
                ("pairs_evaluated", C.c_int64), ("error", C.c_int64), ("ticket", C.c_int64)]


RESULT_WORDS = 12                     # int64 words of mtd_day_result
ROW12 = np.dtype([("a1", "<i4"), ("a2", "<i4"), ("start", "u1"), ("end", "u1"), ("simcelltype", "u1"), ("pad", "u1")])


class TraceTape(C.Structure):
    _fields_ = [("n_scan", C.c_int64), ("scan_key", C.c_void_p), ("scan_u", C.c_void_p),
                ("n_prim", C.c_int64), ("prim_id", C.c_void_p), ("prim_off", C.c_void_p), ("prim_draws", C.c_void_p)]


class TraceStats(C.Structure):
    _fields_ = [("n_primary", C.c_int64), ("n_secondary", C.c_int64), ("n_scan_draws", C.c_int64),
                ("tape_miss", C.c_int64)]


_lib = None


def load():
    """Load the shared library; raises if it has not been built (python -m mtdcovabm_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            raise RuntimeError("libmtd_b200.so is missing: build it with `python mtdcovabm_b200/build.py` "
                               "(there is no CPU fallback for this path)")
        lib = C.CDLL(SO)
        lib.mtd_last_error.restype = C.c_char_p
        lib.mtd_last_error.argtypes = [C.c_void_p]
        for name in EXPORTS:
            getattr(lib, name)            # AttributeError if a declared symbol is not exported
        _lib = lib
    return _lib


def _ptr(a):
    """numpy array / torch tensor / int / None -> c_void_p"""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return C.c_void_p(a.ctypes.data)
    if isinstance(a, int):
        return C.c_void_p(a)
    return C.c_void_p(a.data_ptr())       # torch tensor


class Handle:
    """One device + one stream (SURVEY.md section 8b threading contract)."""

    def __init__(self, n_agents, n_cells, max_intervals, device=0, stream=None, max_rows=0, max_exposures=0):
        self.lib = load()
        cfg = Config(device, n_agents, n_cells, max_intervals, max_rows, max_exposures, stream)
        self.h = C.c_void_p()
        self.n_agents, self.n_cells, self.max_intervals = int(n_agents), int(n_cells), int(max_intervals)
        rc = self.lib.mtd_create(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            msg = self.lib.mtd_last_error(self.h).decode() if self.h else "mtd_create failed"
            if self.h:
                self.lib.mtd_destroy(self.h)
                self.h = None
            raise RuntimeError("mtd_create: rc=%d %s" % (rc, msg))

    def _ck(self, rc, what):
        if rc != 0:
            raise RuntimeError("%s: rc=%d %s" % (what, rc, self.lib.mtd_last_error(self.h).decode()))

    def close(self):
        if getattr(self, "h", None):
            self.lib.mtd_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- setup
    def set_params(self, P: ParamsStruct):
        self._keep_params = P
        self._ck(self.lib.mtd_set_params(self.h, C.byref(P)), "mtd_set_params")

    def set_cells(self, cell_type, cell_indid, mem=MEM_HOST):
        self._ck(self.lib.mtd_set_cells(self.h, C.c_int64(self.n_cells), _ptr(cell_type), _ptr(cell_indid), mem),
                 "mtd_set_cells")

    def set_static(self, age_bracket, epi_age_bracket, soc_rate, res_cellid, mem=MEM_HOST):
        self._ck(self.lib.mtd_set_static(self.h, C.c_int64(self.n_agents), _ptr(age_bracket), _ptr(epi_age_bracket),
                                         _ptr(soc_rate), _ptr(res_cellid), mem), "mtd_set_static")

    def set_state(self, seir=None, inf_type=None, inf_sev=None, last_dose_day=None, sched=None, sched_cnt=None,
                  test_day=None, test_result_day=None, quarantine=None, hosp=None, mem=MEM_HOST):
        self._ck(self.lib.mtd_set_state(self.h, C.c_int64(self.n_agents), _ptr(seir), _ptr(inf_type), _ptr(inf_sev),
                                        _ptr(last_dose_day), _ptr(sched), _ptr(sched_cnt), _ptr(test_day),
                                        _ptr(test_result_day), _ptr(quarantine), _ptr(hosp), mem), "mtd_set_state")

    def get_state(self):
        n = self.n_agents
        out = dict(seir=np.zeros(n, np.int8), inf_type=np.zeros(n, np.int8), inf_sev=np.zeros(n, np.int8),
                   sched=np.zeros((n, MAXSCHED), np.int32), sched_cnt=np.zeros(n, np.uint8),
                   test_day=np.zeros((n, 2), np.int32), test_result_day=np.zeros((n, 2), np.int32),
                   quarantine=np.zeros((n, 3), np.int32), hosp=np.zeros((n, 3), np.int32))
        self._ck(self.lib.mtd_get_state(self.h, C.c_int64(n), _ptr(out["seir"]), _ptr(out["inf_type"]),
                                        _ptr(out["inf_sev"]), _ptr(out["sched"]), _ptr(out["sched_cnt"]),
                                        _ptr(out["test_day"]), _ptr(out["test_result_day"]), _ptr(out["quarantine"]),
                                        _ptr(out["hosp"]), MEM_HOST), "mtd_get_state")
        return out

    def set_tape(self, sel_key, sel_off, sel_picks, pair_k1, pair_k2, pair_u, expo_agent, expo_row):
        t = Tape()
        arrs = [np.ascontiguousarray(sel_key, np.uint64), np.ascontiguousarray(sel_off, np.int32),
                np.ascontiguousarray(sel_picks, np.int32), np.ascontiguousarray(pair_k1, np.uint64),
                np.ascontiguousarray(pair_k2, np.uint32), np.ascontiguousarray(pair_u, np.float64),
                np.ascontiguousarray(expo_agent, np.int32), np.ascontiguousarray(expo_row, np.float64)]
        t.n_sel, t.n_pair, t.n_expo = len(arrs[0]), len(arrs[3]), len(arrs[6])
        (t.sel_key, t.sel_off, t.sel_picks, t.pair_k1, t.pair_k2, t.pair_u, t.expo_agent, t.expo_row) = \
            [a.ctypes.data for a in arrs]
        self._ck(self.lib.mtd_set_tape(self.h, C.byref(t)), "mtd_set_tape")

    def clear_tape(self):
        self._ck(self.lib.mtd_set_tape(self.h, None), "mtd_set_tape")

    # ---- day step
    @staticmethod
    def make_day(mode, day, mhd, testing_enabled=False, quarantine_enabled=False, seed=0):
        return Day(mode, seed, day, mhd, int(testing_enabled), int(quarantine_enabled))

    def simulate_contact_network(self, day: Day, iv_agent, iv_cell, iv_start, iv_end, iv_order=None, n=None,
                                 mem=MEM_HOST):
        n = len(iv_agent) if n is None else n
        self._ck(self.lib.mtd_simulate_contact_network(self.h, C.byref(day), C.c_int64(n), _ptr(iv_agent), _ptr(iv_cell),
                                                       _ptr(iv_start), _ptr(iv_end), _ptr(iv_order), mem),
                 "mtd_simulate_contact_network")

    def simulate_direct_contacts(self, day: Day, delta_buf=None, cap_records=0, apply=True):
        self._ck(self.lib.mtd_simulate_direct_contacts(self.h, C.byref(day), _ptr(delta_buf), C.c_int64(cap_records),
                                                       int(apply)), "mtd_simulate_direct_contacts")

    def set_contacts(self, day: Day, cellid, a1, a2, ov_off, ov_se):
        """the caller's contact pairs of one cell + overlap lists become the day's rows (host numpy arrays)"""
        self._ck(self.lib.mtd_set_contacts(self.h, C.byref(day), C.c_int32(cellid), C.c_int64(len(a1)), _ptr(a1), _ptr(a2),
                                           _ptr(ov_off), _ptr(ov_se), MEM_HOST), "mtd_set_contacts")

    def seed_exposures(self, day: Day, agent_ids, exposure_day=None, pre_exposure_days=0, timesteps=None, asym_mult=None,
                       draws=None):
        """K9: trajectories of agents exposed before the start of the simulation (itinerary.py:373-395 ->
        epidemiology.py:233-500 with reset_states).  Host numpy arrays.  -> int32 [n, 8]: seir, infection type, severity,
        recovered, entries appended, exposure timestep, exposure day, symptomatic day"""
        ids = np.ascontiguousarray(agent_ids, np.int32)
        n = len(ids)
        ts = np.ascontiguousarray(np.arange(144) if timesteps is None else timesteps, np.uint8)
        ed = None if exposure_day is None else np.ascontiguousarray(exposure_day, np.int32)
        am = None if asym_mult is None else np.ascontiguousarray(asym_mult, np.float64)
        dr = None if draws is None else np.ascontiguousarray(draws, np.float64)
        if (ed is not None and len(ed) != n) or (am is not None and len(am) != n) or (dr is not None and dr.size != n * EXPO_W):
            raise ValueError("seed_exposures: per-case arrays must have one entry (draws: EXPO_W entries) per agent id")
        out = np.zeros((n, 8), np.int32)
        self._ck(self.lib.mtd_seed_exposures(self.h, C.byref(day), C.c_int64(n), _ptr(ids), _ptr(ed), C.c_int32(pre_exposure_days),
                                             _ptr(ts), C.c_int32(len(ts)), _ptr(am), _ptr(dr), _ptr(out)), "mtd_seed_exposures")
        return out

    def apply_deltas(self, bufs, n_bufs, cap_records):
        self._ck(self.lib.mtd_apply_deltas(self.h, _ptr(bufs), C.c_int64(n_bufs), C.c_int64(cap_records)),
                 "mtd_apply_deltas")

    # ---- multi-GPU exchange over NVLink peer memory
    def exchange_init(self, rank, world, cap_records) -> bytes:
        """allocate this rank's exchange buffer -> its CUDA IPC handle (bytes) for the peers"""
        buf = C.create_string_buffer(IPC_HANDLE_BYTES)
        self._ck(self.lib.mtd_exchange_init(self.h, int(rank), int(world), C.c_int64(cap_records), buf), "mtd_exchange_init")
        return bytes(buf.raw)

    def exchange_open(self, peer, ipc_handle: bytes):
        self._ck(self.lib.mtd_exchange_open(self.h, int(peer), C.c_char_p(ipc_handle)), "mtd_exchange_open")

    def exchange_ready(self):
        self._ck(self.lib.mtd_exchange_ready(self.h), "mtd_exchange_ready")

    # ---- K8 interval builder
    def build_intervals(self, ent_off, ent_ts, ent_cell, grp_off, grp_agent, grp_rescell, mem=MEM_HOST):
        """itinerary.py:1403-1488 on the device -> length of the built interval stream (use mem=MEM_BUILT next)"""
        n = C.c_int64(0)
        self._ck(self.lib.mtd_build_intervals(self.h, C.c_int64(len(ent_off) - 1), _ptr(ent_off), _ptr(ent_ts), _ptr(ent_cell),
                                              _ptr(grp_off), _ptr(grp_agent), _ptr(grp_rescell), mem, C.byref(n)),
                 "mtd_build_intervals")
        return int(n.value)

    def get_intervals(self, n):
        a, c = np.zeros(n, np.int32), np.zeros(n, np.int32)
        s, e = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        self._ck(self.lib.mtd_get_intervals(self.h, C.c_int64(n), _ptr(a), _ptr(c), _ptr(s), _ptr(e), MEM_HOST), "mtd_get_intervals")
        return a, c, s, e

    # ---- K7 vaccination
    def enable_vaccination(self, n_locals):
        self._ck(self.lib.mtd_enable_vaccination(self.h, C.c_int64(n_locals)), "mtd_enable_vaccination")

    def schedule_vaccinations(self, day: Day, k, replay_ids=None, replay_ts=None):
        """epidemiology.py:911-976 (device part) -> ascending ids of the agents scheduled for day + 1"""
        sel = np.zeros(self.n_agents, np.uint8)
        n = C.c_int64(0)
        nr = 0 if replay_ids is None else len(replay_ids)
        if nr:
            replay_ids = np.ascontiguousarray(replay_ids, np.int32); replay_ts = np.ascontiguousarray(replay_ts, np.int32)
        self._ck(self.lib.mtd_schedule_vaccinations(self.h, C.byref(day), C.c_int64(k), C.c_int64(nr), _ptr(replay_ids),
                                                    _ptr(replay_ts), _ptr(sel), C.byref(n), MEM_HOST),
                 "mtd_schedule_vaccinations")
        ids = np.nonzero(sel)[0].astype(np.int32)
        assert len(ids) == n.value or nr
        return ids

    def set_vaccination(self, last_dose_day=None, vacc_day=None):
        self._ck(self.lib.mtd_set_vaccination(self.h, C.c_int64(self.n_agents), _ptr(last_dose_day), _ptr(vacc_day), MEM_HOST),
                 "mtd_set_vaccination")

    def get_vaccination(self, with_records=True):
        ld = np.zeros(self.n_agents, np.int32)
        vd = np.zeros((self.n_agents, 2), np.int32) if with_records else None
        self._ck(self.lib.mtd_get_vaccination(self.h, C.c_int64(self.n_agents), _ptr(ld), _ptr(vd), MEM_HOST),
                 "mtd_get_vaccination")
        return ld, vd

    def set_intervention_mask(self, on=True):
        """K0 keeps deceased / quarantined / hospitalised agents in their residence cell only"""
        self._ck(self.lib.mtd_set_intervention_mask(self.h, int(on)), "mtd_set_intervention_mask")

    def set_k0_overlap(self, on=True):
        """MEM_DEVICE interval arrays are complete at call time: K0 of the next day may run beside this day's K1..K3"""
        self._ck(self.lib.mtd_set_k0_overlap(self.h, int(on)), "mtd_set_k0_overlap")

    def advance_seir(self, day, mask=None, mem=MEM_HOST):
        self._ck(self.lib.mtd_advance_seir(self.h, C.c_int32(day), _ptr(mask), mem), "mtd_advance_seir")

    def sync(self):
        self._ck(self.lib.mtd_sync(self.h), "mtd_sync")

    # ---- pipelined boundary
    def set_export(self, on=True):
        self._ck(self.lib.mtd_set_export(self.h, int(on)), "mtd_set_export")

    def prefetch_intervals(self, iv_agent, iv_cell, iv_start, iv_end, iv_order=None, n=None):
        """enqueue the upload of the NEXT day's stream (pinned host buffers) on the copy stream"""
        n = len(iv_agent) if n is None else n
        self._ck(self.lib.mtd_prefetch_intervals(self.h, C.c_int64(n), _ptr(iv_agent), _ptr(iv_cell), _ptr(iv_start),
                                                 _ptr(iv_end), _ptr(iv_order)), "mtd_prefetch_intervals")

    def export_day_async(self, rows12, cap_rows, exposed_kv, cap_exposed, result, mem=MEM_HOST):
        """enqueue the export of the day just computed into caller buffers (pinned host or device); -> ticket"""
        tk = C.c_int64(0)
        self._ck(self.lib.mtd_export_day_async(self.h, _ptr(rows12), C.c_int64(cap_rows), _ptr(exposed_kv),
                                               C.c_int64(cap_exposed), _ptr(result), mem, C.byref(tk)),
                 "mtd_export_day_async")
        return int(tk.value)

    def export_wait(self, ticket):
        self._ck(self.lib.mtd_export_wait(self.h, C.c_int64(ticket)), "mtd_export_wait")

    # ---- K5 contact tracing
    def set_residences(self, res_off, res_nres, res_members, mem=MEM_HOST):
        self._ck(self.lib.mtd_set_residences(self.h, C.c_int64(len(res_nres)), _ptr(res_off), _ptr(res_nres),
                                             _ptr(res_members), mem), "mtd_set_residences")

    def contact_tracing(self, day: Day, traced, slice_off, rows3, epi_present=None, tape=None, mem=MEM_HOST):
        """tape: object with scan_key (sorted), scan_u, prim_id, prim_off, prim_draws (numpy) or None (Philox)"""
        t = None
        if tape is not None:
            t = TraceTape()
            keep = [np.ascontiguousarray(tape.scan_key, np.uint64), np.ascontiguousarray(tape.scan_u, np.float64),
                    np.ascontiguousarray(tape.prim_id, np.int32), np.ascontiguousarray(tape.prim_off, np.int32),
                    np.ascontiguousarray(tape.prim_draws, np.float64)]
            t.n_scan, t.n_prim = len(keep[0]), len(keep[2])
            t.scan_key, t.scan_u, t.prim_id, t.prim_off, t.prim_draws = [a.ctypes.data for a in keep]
        slice_off = np.ascontiguousarray(slice_off, np.int64)
        self._ck(self.lib.mtd_contact_tracing(self.h, C.byref(day), C.byref(t) if t is not None else None,
                                              C.c_int64(len(traced)), _ptr(traced), C.c_int64(len(slice_off) - 1),
                                              _ptr(slice_off), _ptr(rows3), _ptr(epi_present), mem),
                 "mtd_contact_tracing")

    def contact_tracing_scan(self, day: Day, traced, slice_off, rows3, primary_out, epi_present=None):
        """cell-sharded K5, phase 1 (DEVICE tensors): this rank's rows -> primary flags into primary_out (u8[N])"""
        slice_off = np.ascontiguousarray(slice_off, np.int64)
        self._ck(self.lib.mtd_contact_tracing_scan(self.h, C.byref(day), C.c_int64(len(traced)), _ptr(traced),
                                                   C.c_int64(len(slice_off) - 1), _ptr(slice_off), _ptr(rows3),
                                                   _ptr(epi_present), _ptr(primary_out), MEM_DEVICE), "mtd_contact_tracing_scan")

    def contact_tracing_finish(self, primary_in):
        """cell-sharded K5, phase 2: the flags merged over the ranks (device u8[N]) -> scheduling on the replicated records"""
        self._ck(self.lib.mtd_contact_tracing_finish(self.h, _ptr(primary_in)), "mtd_contact_tracing_finish")

    def set_k6_tape(self, agent_ids, draws2):
        """replay tape of K6: strictly ascending agent ids, [n, 2] doubles (result draw, quarantine draw)"""
        agent_ids = np.ascontiguousarray(agent_ids, np.int32); draws2 = np.ascontiguousarray(draws2, np.float64)
        assert draws2.shape == (len(agent_ids), 2)
        self._ck(self.lib.mtd_set_k6_tape(self.h, C.c_int64(len(agent_ids)), _ptr(agent_ids), _ptr(draws2)), "mtd_set_k6_tape")

    def process_test_results(self, day: Day, epi_present=None):
        """K6 (itinerary.py:2283-2528, records) -> ascending ids of the agents added to contact_tracing_agent_ids today"""
        ids = np.zeros(self.n_agents, np.int32)
        n = C.c_int64(0)
        if epi_present is not None:
            epi_present = np.ascontiguousarray(epi_present, np.uint8)
        self._ck(self.lib.mtd_process_test_results(self.h, C.byref(day), _ptr(epi_present), _ptr(ids), C.c_int64(len(ids)),
                                                   C.byref(n), MEM_HOST), "mtd_process_test_results")
        return ids[:n.value].copy()

    def process_test_results_device(self, day: Day, ids_dev, epi_present_dev=None):
        """K6 with DEVICE buffers (torch tensors): fills ids_dev (int32) -> number of traced agents"""
        n = C.c_int64(0)
        self._ck(self.lib.mtd_process_test_results(self.h, C.byref(day), _ptr(epi_present_dev), _ptr(ids_dev),
                                                   C.c_int64(ids_dev.numel()), C.byref(n), MEM_DEVICE),
                 "mtd_process_test_results")
        return int(n.value)

    def trace_stats(self) -> "TraceStats":
        st = TraceStats()
        self._ck(self.lib.mtd_get_traced(self.h, None, None, C.byref(st), MEM_HOST), "mtd_get_traced")
        return st

    def schedule_vaccinations_count(self, day: Day, k):
        """K7 without copying the selection flags out -> number of agents scheduled"""
        n = C.c_int64(0)
        self._ck(self.lib.mtd_schedule_vaccinations(self.h, C.byref(day), C.c_int64(k), C.c_int64(0), None, None, None,
                                                    C.byref(n), MEM_HOST), "mtd_schedule_vaccinations")
        return int(n.value)

    def get_traced(self):
        """-> (primary flags uint8[N], secondary flags uint8[N], TraceStats)"""
        prim = np.zeros(self.n_agents, np.uint8); sec = np.zeros(self.n_agents, np.uint8)
        st = TraceStats()
        self._ck(self.lib.mtd_get_traced(self.h, _ptr(prim), _ptr(sec), C.byref(st), MEM_HOST), "mtd_get_traced")
        return prim, sec, st

    # ---- measurement hooks
    def set_profiling(self, on=True):
        self._ck(self.lib.mtd_set_profiling(self.h, int(on)), "mtd_set_profiling")

    def debug_log(self, x):
        """CUDA's fp64 log of every element (test hook)"""
        x = np.ascontiguousarray(x, np.float64)
        out = np.empty_like(x)
        self._ck(self.lib.mtd_debug_log(self.h, C.c_int64(len(x)), _ptr(x), _ptr(out)), "mtd_debug_log")
        return out

    def stage_ms(self):
        """-> ({stage: mean ms over its executions since the previous call}, {stage: kernels launched since then})"""
        ms = (C.c_float * len(STAGES))()
        ln = (C.c_int64 * len(STAGES))()
        self._ck(self.lib.mtd_get_stage_ms(self.h, ms, ln), "mtd_get_stage_ms")
        return dict(zip(STAGES, [float(x) for x in ms])), dict(zip(STAGES, [int(x) for x in ln]))

    # ---- results
    def counters(self) -> Counters:
        c = Counters()
        self._ck(self.lib.mtd_get_counters(self.h, C.byref(c)), "mtd_get_counters")
        return c

    def get_contacts_device(self, rows5, row_dur=None, row_cell=None):
        """export the day's rows into caller-owned DEVICE tensors (int32 [cap,5], [cap], [cap]); -> row count"""
        got = C.c_int64()
        self._ck(self.lib.mtd_get_contacts(self.h, _ptr(rows5), _ptr(row_dur), _ptr(row_cell),
                                           C.c_int64(rows5.shape[0]), C.byref(got), MEM_DEVICE), "mtd_get_contacts")
        return int(got.value)

    def get_contacts(self):
        c = self.counters()
        n = int(c.n_rows)
        rows = np.zeros((max(n, 1), 5), np.int32); dur = np.zeros(max(n, 1), np.int32); cell = np.zeros(max(n, 1), np.int32)
        got = C.c_int64()
        self._ck(self.lib.mtd_get_contacts(self.h, _ptr(rows), _ptr(dur), _ptr(cell), C.c_int64(max(n, 1)), C.byref(got),
                                           MEM_HOST), "mtd_get_contacts")
        return rows[:n], dur[:n], cell[:n]

    def get_exposed(self):
        c = self.counters()
        n = int(c.n_exposed)
        ids = np.zeros(max(n, 1), np.int32)
        got = C.c_int64()
        self._ck(self.lib.mtd_get_exposed(self.h, _ptr(ids), C.c_int64(max(n, 1)), C.byref(got), MEM_HOST),
                 "mtd_get_exposed")
        return ids[:got.value]

    def enable_infection_prob(self):
        self._ck(self.lib.mtd_get_infection_prob(self.h, None, C.c_int64(0), MEM_HOST), "mtd_get_infection_prob")

    def get_infection_prob(self):
        n = int(self.counters().n_rows)
        p = np.zeros(max(n, 1), np.float64)
        self._ck(self.lib.mtd_get_infection_prob(self.h, _ptr(p), C.c_int64(max(n, 1)), MEM_HOST),
                 "mtd_get_infection_prob")
        return p[:n]

    def occupancy(self):
        out = (C.c_int64 * 6)()
        self._ck(self.lib.mtd_occupancy(self.h, out), "mtd_occupancy")
        return np.array(list(out), np.int64)
