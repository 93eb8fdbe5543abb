"""Philox-mode epidemic curves on the reference's recorded itineraries (test infrastructure).

tests/golden/curves_1k.npz holds, for 32 seeds of the unmodified reference (config 1, 20 days), the interval stream of
every day at the contact-network boundary, the arrivals/departures of agents with their state, and the reference's own
incidence / SEIR-occupancy curves.  `run_seed` re-runs the hot path (SEIR advance -> contact network -> transmission)
over those streams with Philox draws, through either the CPU oracle or the CUDA library, and returns the same curves.
"""
from __future__ import annotations

import os

import numpy as np

from mtdcovabm_b200 import params as prm
from oracle import mtd_oracle as mo

HERE = os.path.dirname(os.path.abspath(__file__))
DAYS = 20
_cache = {}
# fixture -> the epidemiology parameter file its reference runs used (sim.py:443)
FIXTURES = {"curves_1k": "epidemiology.json",                       # config 1: 32 seeds x 20 days
            "curves_10k_normal": "epidemiology_normal_days.json",   # config 2: 1k population tiled x10, 8 seeds x 30 days
            "curves_10k_strict": "epidemiology_strict_days.json"}   # config 5


def load(name="curves_1k"):
    if name not in _cache:
        _cache[name] = np.load(os.path.join(HERE, "golden", name + ".npz"))
    return _cache[name]


class SeedData:
    def __init__(self, z, seed):
        p = "s%d_" % seed
        self.N = int(z["n_agents"][0])
        self.days = int(z["days"][0]) if "days" in z.files else DAYS
        self.cell_type = z[p + "cell_type"]; self.cell_indid = z[p + "cell_indid"]
        self.iv_off = z[p + "iv_off"]
        self.iv_cell = z[p + "iv_cell"].astype(np.int32); self.iv_agent = z[p + "iv_agent"].astype(np.int32)
        self.iv_start = z[p + "iv_start"]; self.iv_end = z[p + "iv_end"]
        self.present = np.unpackbits(z[p + "present"], axis=1)[:, :self.N].astype(bool)
        self.arr_state, self.arr_sched, self.last_dose = z[p + "arr_state"], z[p + "arr_sched"], z[p + "last_dose"]
        self.dyn = z[p + "dyn"]
        self.ref_incidence, self.ref_occupancy = z[p + "ref_incidence"], z[p + "ref_occupancy"]
        self.static = {k: z[p + k] for k in ("age_bracket", "epi_age_bracket", "soc_rate", "res_cellid")}

    def initial_state(self):
        st = mo.AgentState(self.N)
        for k, v in self.static.items():
            getattr(st, k)[:] = v
        return st

    def day_stream(self, i):
        a, b = self.iv_off[i], self.iv_off[i + 1]
        return (np.ascontiguousarray(self.iv_cell[a:b]), np.ascontiguousarray(self.iv_agent[a:b]),
                np.ascontiguousarray(self.iv_start[a:b]), np.ascontiguousarray(self.iv_end[a:b]))

    def apply_arrivals(self, st, i):
        """agents that appear on day index i get the state the reference gave them; departed agents lose theirs"""
        day = i + 1
        if i > 0:
            gone = self.present[i - 1] & ~self.present[i]
            st.seir[gone] = 0; st.inf_type[gone] = 0; st.inf_sev[gone] = 0; st.sched_cnt[gone] = 0
        for (d, a, s, t, v) in self.arr_state[self.arr_state[:, 0] == day]:
            st.seir[a] = s; st.inf_type[a] = t; st.inf_sev[a] = v; st.sched_cnt[a] = 0
            st.test_day[a] = st.test_result_day[a] = mo.NONE
            st.quarantine[a] = st.hosp[a] = mo.NONE
        for (d, a, dd, t, ts) in self.arr_sched[self.arr_sched[:, 0] == day]:
            st.sched[a, st.sched_cnt[a]] = mo.pack_sched(dd, t, ts); st.sched_cnt[a] += 1
        for (d, a, v) in self.last_dose[self.last_dose[:, 0] == day]:
            st.last_dose_day[a] = v


def run_seed_oracle(sd: SeedData, P, philox_seed):
    """-> (incidence[20], occupancy[20, 6]) of the CPU oracle in Philox mode"""
    st = sd.initial_state()
    inc = np.zeros(sd.days, np.int32); occ = np.zeros((sd.days, 6), np.int32)
    for i in range(sd.days):
        day = i + 1
        mo.advance(day, st, None)
        sd.apply_arrivals(st, i)
        cell, agent, s, e = sd.day_stream(i)
        mhd, te, qe = sd.dyn[i]
        ctx = mo.make_ctx(mo.MODE_PHILOX, day, float(mhd), bool(te), bool(qe), seed=philox_seed)
        rows, dur, rc, _ = mo.contact_network(P, ctx, cell, agent, s, e, sd.cell_type, sd.cell_indid, st)
        exp, _, _ = mo.transmission(P, ctx, cell, agent, s, e, sd.cell_type, rows, dur, rc, st)
        inc[i] = len(exp)
        occ[i] = np.bincount(st.seir[sd.present[i]].astype(np.int64), minlength=6)[:6]
    return inc, occ


def run_seed_gpu(sd: SeedData, P, philox_seed):
    """same day loop through libmtd_b200.so (the state lives on the device; arrivals are patched in from the host)"""
    from mtdcovabm_b200 import capi
    from tests import helpers as H
    st = sd.initial_state()
    h = H.make_handle(P, st, sd.cell_type, sd.cell_indid, int(np.diff(sd.iv_off).max()) + 16)
    inc = np.zeros(sd.days, np.int32); occ = np.zeros((sd.days, 6), np.int32)
    for i in range(sd.days):
        day = i + 1
        h.advance_seir(day, None)
        dev = h.get_state()
        for f in H.STATE_FIELDS:
            getattr(st, f)[...] = dev[f]
        sd.apply_arrivals(st, i)
        H.upload_state(h, st)
        cell, agent, s, e = sd.day_stream(i)
        mhd, te, qe = sd.dyn[i]
        d = h.make_day(capi.MODE_PHILOX, day, float(mhd), bool(te), bool(qe), seed=philox_seed)
        h.simulate_contact_network(d, agent, cell, s, e)
        h.simulate_direct_contacts(d)
        inc[i] = len(h.get_exposed())
        seir = h.get_state()["seir"]
        occ[i] = np.bincount(seir[sd.present[i]].astype(np.int64), minlength=6)[:6]
    h.close()
    return inc, occ


def ks_table(ours_inc, ours_occ, ref_inc, ref_occ):
    """two-sample KS p-value per day for incidence and each occupied SEIR state -> dict name -> p[20]"""
    from scipy.stats import ks_2samp
    days = ours_inc.shape[1]
    out = {"incidence": np.array([ks_2samp(ours_inc[:, d], ref_inc[:, d]).pvalue for d in range(days)])}
    for k, name in ((1, "S"), (2, "E"), (3, "I"), (4, "R"), (5, "D")):
        out[name] = np.array([ks_2samp(ours_occ[:, d, k], ref_occ[:, d, k]).pvalue for d in range(days)])
    return out


def params_config1():
    return prm.build_params(None, prm.load_json("epidemiology.json"))


def params_for(name):
    return prm.build_params(None, prm.load_json(FIXTURES[name]))
