"""Bring-up + instrumentation of the *unmodified* reference (jcus0006/mtdcovabm, /root/reference).

TEST INFRASTRUCTURE ONLY.  This module runs in the build container (where /root/reference exists) to
 (a) import the reference's pure-Python simulator with stubs for absent third-party modules that the hot
     path never calls, and
 (b) record, at the `ContactNetwork.simulate_contact_network` boundary (contactnetwork.py:65), per-day
     snapshots (inputs, outputs) and a *keyed draw tape* of every random draw the hot path consumed
     (SURVEY.md Appendix C).
The recorded fixtures are written by make_golden.py as .npz files under tests/golden/ and are what pins the
C oracle (oracle/) and the CUDA path.  Nothing here is imported by the product package.
"""
from __future__ import annotations

import os
import random
import shutil
import sys
import types

import numpy as np

REF_ROOT = "/root/reference"

# draw-tape "exposure record" layout (one row of float64 per exposed agent, NaN = draw not consumed)
# [0] exposure timestep (epidemiology.py:246)
# [1..10] rounded log-normal day offsets indexed by SEIRStateTransition id 0..9 (epidemiology.py:248..450)
# [11] symptomatic u (:268)  [12] severe u (:295)  [13] critical u (:319)  [14] deceased u (:341)
# [15] test timestep (:480)  [16] test-after-symptoms u (:527)  [17] testing days (:537)  [18] result days (:546)
# [19] quarantine timestep (:496) [20] quarantine u (:575)
EXPO_W = 21


def _install_stubs():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    class _Dummy:
        def __init__(self, *a, **k):
            pass

        def __getattr__(self, k):
            return _Dummy()

        def __call__(self, *a, **k):
            return _Dummy()

        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

    class _Permissive(types.ModuleType):
        def __getattr__(self, k):
            if k.startswith("__"):
                raise AttributeError(k)
            return _Dummy

    if "matplotlib" not in sys.modules:
        try:
            import matplotlib  # noqa: F401
        except Exception:
            mpl = mod("matplotlib")
            mpl.pyplot = mod("matplotlib.pyplot")
    mod("pympler", asizeof=mod("pympler.asizeof", asizeof=lambda x: 0))
    sys.modules["pympler"].asizeof = sys.modules["pympler.asizeof"]

    def profile(func=None, **k):
        if func is None:
            return lambda f: f
        return func

    mod("memory_profiler", profile=profile)
    mod("bintrees", FastAVLTree=dict)

    class Power_Law:
        def __init__(self, xmin=1.0, parameters=None, **k):
            self.xmin = xmin
            self.alpha = parameters[0]

        def generate_random(self, n):
            u = np.random.random(n)
            return self.xmin * (1.0 - u) ** (-1.0 / (self.alpha - 1.0))

    mod("powerlaw", Power_Law=Power_Law)
    for name in ("dask", "dask.distributed", "dask.dataframe", "dask.array", "dask.bag", "distributed",
                 "distributed.diagnostics", "distributed.diagnostics.plugin", "dask.config"):
        sys.modules[name] = _Permissive(name)
    sys.modules["dask"].distributed = sys.modules["dask.distributed"]
    sys.modules["dask"].config = sys.modules["dask.config"]
    sys.modules["distributed"].diagnostics = sys.modules["distributed.diagnostics"]


def prepare_workdir(workdir, popsubfolder="1kagents2ktourists2019", epidemiology_variant=None, pop_src=None):
    """Writable copy of the files sim.py opens relative to the CWD (Appendix F step 1).  pop_src: a population
    directory outside the reference tree (the tiled 10k population of make_pop10k.py)."""
    if os.path.exists(workdir):
        shutil.rmtree(workdir)
    os.makedirs(workdir)
    shutil.copytree(os.path.join(REF_ROOT, "simulator"), os.path.join(workdir, "simulator"))
    shutil.copytree(os.path.join(REF_ROOT, "data"), os.path.join(workdir, "data"))
    shutil.copy(os.path.join(REF_ROOT, "config.json"), os.path.join(workdir, "config.json"))
    os.makedirs(os.path.join(workdir, "population"))
    shutil.copytree(pop_src if pop_src is not None else os.path.join(REF_ROOT, "population", popsubfolder),
                    os.path.join(workdir, "population", popsubfolder))
    if epidemiology_variant is not None:  # sim.py:443 hard-codes data/epidemiology.json
        shutil.copy(os.path.join(REF_ROOT, "data", epidemiology_variant),
                    os.path.join(workdir, "data", "epidemiology.json"))
    return workdir


class Recorder:
    """Snapshots + keyed draw tape, filled by the monkey-patches installed in run_reference()."""

    def __init__(self):
        self.days = []          # list of dict per simulated day
        self.cur = None         # current day record
        self.ctx_cell = None    # cell currently inside generate_contact_network / simulate_direct_contacts
        self.ctx_pair = None    # pair currently evaluated
        self.ctx_expo = None    # exposure row being filled
        self.ctx_site = None
        self.enabled = False
        self.advance_calls = []  # KAT records of seirstateutil.update_agent_state (seirstateutil.py:38)
        self.in_uas = False
        self.ct_active = False   # inside Epidemiology.contact_tracing
        self.ct_events = []      # flat, ordered draw / call events of the current contact_tracing call
        self.traces = []         # one record per contact_tracing call that had work to do
        self.residences = None   # {res_cellid: (resident ids, staff ids)}
        self.ext_ids = set()     # agents whose SEIR state the itinerary stage set since the last boundary
        self.n_agents_total = None


def run_reference(workdir, seed, days, popsubfolder="1kagents2ktourists2019", record_days=None, quiet=True,
                  on_day=None, extra_install=None):
    """Run sim.main() of the unmodified reference and return the Recorder."""
    _install_stubs()
    simdir = os.path.join(workdir, "simulator")
    if simdir not in sys.path:
        sys.path.insert(0, simdir)
    os.chdir(workdir)
    random.seed(seed)
    np.random.seed(seed)

    import sim  # noqa: E402  (reference module)

    sim.params.update({"popsubfolder": popsubfolder, "simulationdays": days, "numprocesses": 1, "use_mp": True,
                       "logmemoryinfo": False, "logfilename": "golden_%d.txt" % seed})
    rec = Recorder()
    install_recorder(rec, record_days, on_day)
    if extra_install is not None:          # further patches of a fixture script (make_k6_kat.py)
        extra_install(rec)
    stdout = sys.stdout
    try:
        sim.main()
    finally:
        try:
            if sys.stdout is not stdout and hasattr(sys.stdout, "close"):
                sys.stdout.close()
        except Exception:
            pass
        sys.stdout = stdout
    return rec


def install_recorder(rec, record_days=None, on_day=None):
    """Monkey-patch the imported reference modules (simulator/ must be on sys.path) so that every call of
    ContactNetwork.simulate_contact_network fills `rec` with the boundary snapshot and the keyed draw tape."""
    import contactnetwork as cn  # noqa: E402
    import epidemiology as epi  # noqa: E402
    import util as rutil  # noqa: E402

    # ---------------- boundary snapshot ----------------
    orig_sim_cn = cn.ContactNetwork.simulate_contact_network

    def flat_epi(agents_epi, ids):
        out = {}
        for i in ids:
            e = agents_epi.get(i) if hasattr(agents_epi, "get") else agents_epi[i]
            if e is None:
                continue
            out[int(i)] = {k: (None if v is None else [[int(y) for y in x] if isinstance(x, (list, tuple, np.ndarray)) else int(x) for x in v])
                           for k, v in e.items() if k in ("state_transition_by_day", "test_day", "test_result_day",
                                                           "hospitalisation_days", "quarantine_days")}
        return out

    def snap_state(self):
        vu = self.vars_util
        ids = sorted(int(k) for k in vu.agents_seir_state.keys())
        return {
            "ids": np.array(ids, dtype=np.int32),
            "seir": np.array([int(vu.agents_seir_state[i]) for i in ids], dtype=np.int8),
            "inf_type": {int(k): int(v) for k, v in vu.agents_infection_type.items()},
            "inf_sev": {int(k): int(v) for k, v in vu.agents_infection_severity.items()},
            "vacc": {int(k): [int(d) for d in v] for k, v in vu.agents_vaccination_doses.items()},
        }

    def patched_sim_cn(self, day, weekday):
        want = record_days is None or day in record_days
        if not want:
            return orig_sim_cn(self, day, weekday)
        vu = self.vars_util
        d = {"day": int(day), "weekday": int(weekday)}
        d["n_agents_total"] = int(self.epi_util.n_locals + self.epi_util.n_tourists)
        d["n_locals"] = int(self.epi_util.n_locals)
        d["all_cell_type"] = {int(k): int(v) for k, v in self.cells_type.items()}
        d["all_cell_indid"] = {int(k): int(v) for k, v in self.indids_by_cellid.items() if v is not None}
        dp = self.epi_util.dyn_params
        d["mhd"] = float(dp.masks_hygiene_distancing_multiplier)
        d["testing_enabled"] = bool(dp.testing_enabled)
        d["quarantine_enabled"] = bool(dp.quarantine_enabled)
        d["contact_tracing_enabled"] = bool(dp.contact_tracing_enabled)
        cells, agents, st, en = [], [], [], []
        for cellid, lst in vu.cells_agents_timesteps.items():
            for (a, s, e) in lst:
                cells.append(int(cellid)); agents.append(int(a)); st.append(int(s)); en.append(int(e))
        d["iv_cell"] = np.array(cells, dtype=np.int32)
        d["iv_agent"] = np.array(agents, dtype=np.int32)
        d["iv_start"] = np.array(st, dtype=np.uint8)
        d["iv_end"] = np.array(en, dtype=np.uint8)
        act = sorted(set(agents))
        d["static_ids"] = np.array(act, dtype=np.int32)
        d["age_bracket"] = np.array([int(self.agents_static.get(a, "age_bracket_index")) for a in act], dtype=np.int8)
        d["epi_age_bracket"] = np.array([int(self.agents_static.get(a, "epi_age_bracket_index")) for a in act], dtype=np.int8)
        d["soc_rate"] = np.array([float(self.agents_static.get(a, "soc_rate")) for a in act], dtype=np.float64)
        d["res_cellid"] = np.array([int(self.agents_static.get(a, "res_cellid")) for a in act], dtype=np.int32)
        ucells = sorted(set(cells))
        d["cell_ids"] = np.array(ucells, dtype=np.int32)
        d["cell_type"] = np.array([int(self.cells_type[c]) for c in ucells], dtype=np.int8)
        d["cell_indid"] = np.array([int(self.indids_by_cellid[c]) if (c in self.indids_by_cellid and self.indids_by_cellid[c] is not None) else -1
                                    for c in ucells], dtype=np.int16)
        d["pre"] = snap_state(self)
        d["ext_ids"] = sorted(rec.ext_ids)   # entry / re-entry / day-1 infections set by itinerary.py:392,645,1578,1584
        d["ext_epi"] = flat_epi(self.agents_epi, d["ext_ids"])
        rec.ext_ids.clear()
        d["pre_epi"] = flat_epi(self.agents_epi, act)
        n_before = len(vu.directcontacts_by_simcelltype_by_day)
        d["tape_sel"] = []      # (cell, main_agent, n_candidates, [indices into unshuffled candidate list])
        d["tape_pair_u"] = []   # (cell, a1, a2, u, p)
        d["tape_expo"] = []     # (cell, a1, a2, agent, row[EXPO_W])
        d["cell_pc"] = []       # debugging aid: per-cell potential-contact count
        rec.cur = d
        rec.enabled = True
        try:
            out = orig_sim_cn(self, day, weekday)
        finally:
            rec.enabled = False
        rows = vu.directcontacts_by_simcelltype_by_day[n_before:]
        d["contacts"] = np.array([[int(x) for x in r] for r in rows], dtype=np.int32).reshape(-1, 5)
        d["exposed_ids"] = np.array([int(x) for x in out[1]], dtype=np.int32)
        d["post"] = snap_state(self)
        d["post_epi"] = flat_epi(self.agents_epi, act)
        rec.days.append(d)
        rec.cur = None
        if on_day is not None:
            on_day(d)
        return out

    cn.ContactNetwork.simulate_contact_network = patched_sim_cn

    # ---------------- SEIR daily advance KAT (seirstateutil.py:38-72) ----------------
    import seirstateutil as ssu  # noqa: E402
    orig_uas = ssu.update_agent_state

    def patched_uas(agents_seir_state, agents_infection_type, agents_infection_severity, agentid, agent, agentindex, day):
        rec.in_uas = True
        try:
            return _patched_uas(agents_seir_state, agents_infection_type, agents_infection_severity, agentid, agent, agentindex, day)
        finally:
            rec.in_uas = False

    def _patched_uas(agents_seir_state, agents_infection_type, agents_infection_severity, agentid, agent, agentindex, day):
        sched = agent["state_transition_by_day"]
        if not sched:
            return orig_uas(agents_seir_state, agents_infection_type, agents_infection_severity, agentid, agent, agentindex, day)
        before = [[int(y) for y in x] for x in sched]
        pre = (int(agents_seir_state[agentid]),
               int(agents_infection_type[agentid]) if agentid in agents_infection_type else 0,
               int(agents_infection_severity[agentid]) if agentid in agents_infection_severity else 0)
        out = orig_uas(agents_seir_state, agents_infection_type, agents_infection_severity, agentid, agent, agentindex, day)
        post = (int(agents_seir_state[agentid]),
                int(agents_infection_type[agentid]) if agentid in agents_infection_type else 0,
                int(agents_infection_severity[agentid]) if agentid in agents_infection_severity else 0)
        after = [[int(y) for y in x] for x in agent["state_transition_by_day"]]
        rec.advance_calls.append((int(day), int(agentid), before, pre, post, after, out is not None))
        return out

    ssu.update_agent_state = patched_uas

    orig_assu = ssu.agents_seir_state_update

    def patched_assu(agents_seir_state, new_seir_state, agent_id_or_idx, *a, **k):
        if not rec.enabled and not rec.in_uas:   # outside the contact-network stage and not the daily advance
            rec.ext_ids.add(int(agent_id_or_idx))
        return orig_assu(agents_seir_state, new_seir_state, agent_id_or_idx, *a, **k)

    ssu.agents_seir_state_update = patched_assu

    # ---------------- per-cell context ----------------
    orig_by_cell = cn.ContactNetwork.simulate_contact_network_by_cellid

    def patched_by_cell(self, cellid, day):
        rec.ctx_cell = int(cellid)
        try:
            return orig_by_cell(self, cellid, day)
        finally:
            rec.ctx_cell = None

    cn.ContactNetwork.simulate_contact_network_by_cellid = patched_by_cell

    # ---------------- selection tape (util.py:274-297 + contactnetwork.py:299) ----------------
    orig_get_all = rutil.get_all_contacts_ids_by_id
    sel_state = {}

    def patched_get_all(id, keys, agents_ids=None, agents_degrees=None, shuffle=False):
        if not (rec.enabled and shuffle):
            return orig_get_all(id, keys, agents_ids, agents_degrees, shuffle)
        unshuffled = orig_get_all(id, keys, agents_ids, agents_degrees, False)
        if unshuffled is None:
            return None
        shuffled = np.array(unshuffled)
        np.random.shuffle(shuffled)  # same single RNG call the reference makes (util.py:295)
        sel_state["main"] = int(id)
        sel_state["unshuffled"] = [int(x) for x in unshuffled]
        sel_state["shuffled"] = shuffled
        return shuffled

    rutil.get_all_contacts_ids_by_id = patched_get_all

    orig_choice = np.random.choice

    def patched_choice(a, size=None, replace=True, p=None):
        res = orig_choice(a, size=size, replace=replace, p=p)
        if rec.ct_active:
            arr = np.asarray(a)
            if arr.shape == (79,) and arr[0] == 42 and size == 1:
                rec.ct_events.append(("ts", int(res[0])))
            else:
                rec.ct_events.append(("sel", [int(x) for x in res]))
        if not rec.enabled:
            return res
        if sel_state.get("shuffled") is not None and a is sel_state["shuffled"]:
            pos = {v: i for i, v in enumerate(sel_state["unshuffled"])}
            rec.cur["tape_sel"].append((rec.ctx_cell, sel_state["main"], len(sel_state["unshuffled"]),
                                        [pos[int(x)] for x in res]))
            sel_state["shuffled"] = None
        elif rec.ctx_expo is not None:
            site = rec.ctx_site
            if site == "expo_ts":
                rec.ctx_expo[0] = float(res[0]); rec.ctx_site = "after_ts"
            elif site == "test_ts":
                rec.ctx_expo[15] = float(res[0])
            elif site == "quar_ts":
                rec.ctx_expo[19] = float(res[0])
        return res

    np.random.choice = patched_choice

    # ---------------- transmission tape ----------------
    orig_random = random.random
    orig_sdc = epi.Epidemiology.simulate_direct_contacts
    orig_trans = epi.Epidemiology.simulate_seir_state_transition
    orig_test = epi.Epidemiology.schedule_test
    orig_quar = epi.Epidemiology.schedule_quarantine
    orig_sln = rutil.sample_log_normal

    class PairTracker(dict):
        """dict wrapper whose .items() announces the pair being evaluated (epidemiology.py:91)."""

        def items(self):
            for k, v in dict.items(self):
                rec.ctx_pair = (int(k[0]), int(k[1]))
                yield k, v
            rec.ctx_pair = None

    def patched_sdc(self, agents_directcontacts, cellid, cell_type, day):
        if not rec.enabled:
            return orig_sdc(self, agents_directcontacts, cellid, cell_type, day)
        return orig_sdc(self, PairTracker(agents_directcontacts), cellid, cell_type, day)

    epi.Epidemiology.simulate_direct_contacts = patched_sdc

    def patched_random():
        u = orig_random()
        if rec.ct_active:
            rec.ct_events.append(("u", u))
        if rec.enabled:
            if rec.ctx_expo is not None:
                site = rec.ctx_site
                idx = {"symp": 11, "sev": 12, "cri": 13, "dec": 14, "test_u": 16, "quar_u": 20}.get(site)
                if idx is None:
                    # branch draws arrive in fixed order: symptomatic, severe, critical, deceased
                    for idx in (11, 12, 13, 14):
                        if np.isnan(rec.ctx_expo[idx]):
                            break
                rec.ctx_expo[idx] = u
            elif rec.ctx_pair is not None:
                rec.cur["tape_pair_u"].append((rec.ctx_cell, rec.ctx_pair[0], rec.ctx_pair[1], u))
        return u

    random.random = patched_random
    epi.random.random = patched_random

    def patched_trans(self, current_day, exposed_agent, exposed_agent_id, incremental_days, overlapping_timesteps,
                      agent_state_transition_by_day, agent_epi_age_bracket_index, agent_quarantine_days,
                      asymptomatic_multiplier=1.0):
        if not (rec.enabled and rec.ctx_pair is not None):
            return orig_trans(self, current_day, exposed_agent, exposed_agent_id, incremental_days,
                              overlapping_timesteps, agent_state_transition_by_day, agent_epi_age_bracket_index,
                              agent_quarantine_days, asymptomatic_multiplier)
        row = np.full(EXPO_W, np.nan)
        rec.ctx_expo = row
        rec.ctx_site = "expo_ts"
        n_before = len(agent_state_transition_by_day)
        try:
            out = orig_trans(self, current_day, exposed_agent, exposed_agent_id, incremental_days,
                             overlapping_timesteps, agent_state_transition_by_day, agent_epi_age_bracket_index,
                             agent_quarantine_days, asymptomatic_multiplier)
        finally:
            rec.ctx_expo = None
            rec.ctx_site = None
        # day offsets per transition id, recovered from the appended schedule entries (cumulative days)
        prev = current_day
        for (dday, trans, ts) in out[0][n_before:]:
            row[1 + int(trans)] = float(int(dday) - prev)
            prev = int(dday)
        rec.cur["tape_expo"].append((rec.ctx_cell, rec.ctx_pair[0], rec.ctx_pair[1], int(exposed_agent_id), row,
                                     [int(t) for t in overlapping_timesteps]))
        return out

    epi.Epidemiology.simulate_seir_state_transition = patched_trans

    def patched_test(self, agent, agent_id, incremental_days, start_timestep, quarantine_type):
        if rec.ct_active:
            rec.ct_events.append(("st", int(agent_id), int(quarantine_type)))
        if rec.ctx_expo is None:
            return orig_test(self, agent, agent_id, incremental_days, start_timestep, quarantine_type)
        rec.ctx_site = "test_u"
        lognormals = []
        rec._ln_sink = lognormals
        try:
            out = orig_test(self, agent, agent_id, incremental_days, start_timestep, quarantine_type)
        finally:
            rec._ln_sink = None
            rec.ctx_site = "quar_ts"
        if len(lognormals) == 2:
            rec.ctx_expo[17], rec.ctx_expo[18] = float(lognormals[0]), float(lognormals[1])
        return out

    epi.Epidemiology.schedule_test = patched_test

    def patched_quar(self, agent_id, start_day, start_timestep, quarantine_type, end_day=None, quarantine_days=None,
                     agent=None):
        if rec.ct_active:
            rec.ct_events.append(("sq", int(agent_id), int(quarantine_type)))
        if rec.ctx_expo is None:
            return orig_quar(self, agent_id, start_day, start_timestep, quarantine_type, end_day, quarantine_days, agent)
        rec.ctx_site = "quar_u"
        try:
            return orig_quar(self, agent_id, start_day, start_timestep, quarantine_type, end_day, quarantine_days, agent)
        finally:
            rec.ctx_site = None

    epi.Epidemiology.schedule_quarantine = patched_quar
    rec._ln_sink = None

    def patched_sln(mean, std, size, isInt=False, recomputeparams=True):
        v = orig_sln(mean, std, size, isInt, recomputeparams)
        if rec.ct_active:
            rec.ct_events.append(("ln", int(v)))
        if rec.enabled and rec._ln_sink is not None:
            rec._ln_sink.append(v)
        return v

    rutil.sample_log_normal = patched_sln

    # the trajectory draws "test_ts" happens right before schedule_test: mark via a wrapper on the site name
    # (np.random.choice(timestep_options) at epidemiology.py:480): after the trajectory the next choice is test ts.
    orig_hosp = epi.Epidemiology.schedule_hospitalisation

    def patched_hosp(self, agent, agent_id, hospitalisation_days):
        return orig_hosp(self, agent, agent_id, hospitalisation_days)

    epi.Epidemiology.schedule_hospitalisation = patched_hosp

    # site sequencing for the two timestep_options draws: first after trajectory -> test_ts, second -> quar_ts
    def site_aware_choice(a, size=None, replace=True, p=None):
        if rec.enabled and rec.ctx_expo is not None and rec.ctx_site == "after_ts" and a is not None \
                and isinstance(a, np.ndarray) and a.shape == (79,) and a[0] == 42:
            rec.ctx_site = "test_ts"
        return patched_choice(a, size=size, replace=replace, p=p)

    np.random.choice = site_aware_choice


    # ---------------- contact tracing (epidemiology.py:667-834) ----------------
    orig_ct = epi.Epidemiology.contact_tracing

    def all_epi(self):
        keys = [int(k) for k in (self.agents_epi.keys() if hasattr(self.agents_epi, "keys") else range(len(self.agents_epi)))]
        return flat_epi(self.agents_epi, keys)

    def patched_ct(self, day, distributed=False, f=None):
        vu = self.vars_util
        traced = sorted(int(x) for x in vu.contact_tracing_agent_ids)
        if not (self.dyn_params.contact_tracing_enabled and len(traced) > 0):
            return orig_ct(self, day, distributed, f)
        if record_days is not None and day not in record_days:     # long runs: only the kept days carry a trace record
            return orig_ct(self, day, distributed, f)
        if rec.residences is None:
            rec.residences = {}
            rec._hh_inst = set()
            for cells, rk, sk in ((self.cells_households, "resident_uids", "staff_uids"),
                                  (self.cells_institutions, "resident_uids", "staff_uids"),
                                  (self.cells_accommodation, "member_uids", None)):
                for cid, cellv in cells.items():
                    if int(cid) in rec.residences:      # households win over institutions over accommodation (:752-763)
                        continue
                    place = cellv["place"]
                    res = [int(x) for x in (place.get(rk) if place.get(rk) is not None else [])]
                    stf = [int(x) for x in (place.get(sk) if sk and place.get(sk) is not None else [])]
                    rec.residences[int(cid)] = (res, stf)
                    if sk is not None:
                        rec._hh_inst.add(int(cid))
        t = {"day": int(day), "traced": traced,
             "testing_enabled": bool(self.dyn_params.testing_enabled),
             "quarantine_enabled": bool(self.dyn_params.quarantine_enabled),
             "rows": np.array([[int(x) for x in r[:3]] for r in vu.directcontacts_by_simcelltype_by_day],
                              dtype=np.int32).reshape(-1, 3),
             "markers": {int(k): int(v) for k, v in vu.directcontacts_by_simcelltype_by_day_start_marker.items()},
             "n_locals": int(self.n_locals)}
        t["pre_epi"] = all_epi(self)
        resc = {}
        for a in t["pre_epi"].keys():
            try:
                resc[a] = int(self.agents_static.get(a, "res_cellid"))
            except Exception:
                resc[a] = None
        t["res_cellid"] = resc
        # accommodation member lists change as tourists come and go: snapshot them at every call
        t["accom"] = {int(cid): [int(x) for x in cellv["place"].get("member_uids", [])]
                      for cid, cellv in self.cells_accommodation.items()}
        rec.ct_events = []
        rec.ct_active = True
        try:
            out = orig_ct(self, day, distributed, f)
        finally:
            rec.ct_active = False
        t["events"] = rec.ct_events
        t["updated"] = [int(x) for x in out[1]]
        t["post_epi"] = all_epi(self)
        rec.traces.append(t)
        return out

    epi.Epidemiology.contact_tracing = patched_ct
