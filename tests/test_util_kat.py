"""Known-answer tests of the hot-path helpers against vectors produced by the unmodified reference
(tests/golden/make_util_kat.py -> util_kat.npz): the host mirror in mtdcovabm_b200/util.py, the C oracle's transition
table, and -- on the GPU -- the K4 kernel's transition table over all 6 x 5 x 4 x 10 inputs."""
import ctypes as C
import os

import numpy as np
import pytest

from mtdcovabm_b200 import util
from mtdcovabm_b200.epidemiologyclasses import CellType
from oracle import mtd_oracle as mo

Z = np.load(os.path.join(os.path.dirname(__file__), "golden", "util_kat.npz"))


def test_sample_log_normal_matches_reference_stream():
    for k, (m, s) in enumerate(Z["ln_pars"]):
        np.random.seed(100 + k)
        got = [util.sample_log_normal(m, s, 1, True) for _ in range(8)]
        assert got == [int(x) for x in Z["ln_int_draws"][k]]
    np.random.seed(7)
    assert np.array_equal(np.asarray(util.sample_log_normal(4.5, 1.5, 16, False)), Z["ln_float_draws"])


def test_lognormal_parameter_conversion_matches_kernel_parameters():
    """the (mu, sigma) the kernels use (params.build_params -> ln_mu / ln_sigma) are the reference's conversion"""
    from mtdcovabm_b200 import params as prm
    P = prm.build_params()
    ep = prm.load_json("epidemiology.json") if hasattr(prm, "load_json") else None
    mu, sigma = util.lognormal_params(4.5, 1.5)
    assert mu == np.log(4.5 ** 2 / np.sqrt(1.5 ** 2 + 4.5 ** 2)) and sigma == np.sqrt(np.log(1.5 ** 2 / 4.5 ** 2 + 1))
    assert all(np.isfinite(P.ln_mu[i]) and P.ln_sigma[i] > 0 for i in range(10))


def test_vaccination_multipliers_kat():
    for d in range(0, 401):
        i, a, acc = util.calculate_vaccination_multipliers({5: [10]}, 5, 10 + d, 0.9, 0.8, 30)
        assert (i, a) == tuple(Z["vacc_mult"][d]) and acc is None
    assert util.calculate_vaccination_multipliers({5: [10]}, 6, 50, 0.9, 0.8, 30)[:2] == tuple(Z["vacc_none"])
    assert util.calculate_vaccination_multipliers({}, 1, 3, 0.9, 0.8, 30, accum_time=0.0)[2] >= 0.0


def test_simcelltype_table_kat():
    got = [int(util.convert_celltype_to_simcelltype(0, celltype=CellType(t))) for t in range(11)]
    assert got == [int(x) for x in Z["simcelltype"]]
    assert int(util.convert_celltype_to_simcelltype(3, cells={3: {"type": CellType.Classroom}})) == 2


def test_partner_lists_kat():
    keys = [(1, 2), (3, 1), (2, 3), (4, 1), (2, 4), (5, 6)]
    ids, deg = [1, 2, 3, 4, 5, 6], [2, 0, 1, 3, 1, 0]
    assert list(util.get_all_contacts_ids_by_id(1, keys)) == [int(x) for x in Z["partners_plain"][0]]
    assert list(util.get_all_contacts_ids_by_id(1, keys, ids, deg)) == [int(x) for x in Z["partners_alive"]]
    np.random.seed(11)
    assert list(util.get_all_contacts_ids_by_id(1, keys, ids, [1] * 6, shuffle=True)) == [int(x) for x in Z["partners_shuffled"]]
    assert bool(Z["partners_none"][0]) and util.get_all_contacts_ids_by_id(5, keys, ids, deg) is None


def test_oracle_transition_table_matches_reference_on_every_input():
    L = mo.lib()
    ns, nt, nv = C.c_int(), C.c_int(), C.c_int()
    for s, it, sv, tr, es, et, ev in Z["transition_table"]:
        L.mtdo_transition(int(s), int(it), int(sv), int(tr), C.byref(ns), C.byref(nt), C.byref(nv))
        assert (ns.value, nt.value, nv.value) == (es, et, ev), (s, it, sv, tr)


@pytest.mark.gpu
def test_k4_transition_table_matches_reference_on_every_input():
    """one agent per (state, type, severity, transition) with a single entry dated today -> K4 -> the reference's table"""
    from mtdcovabm_b200 import capi, params as prm
    T = Z["transition_table"]
    n = len(T)
    h = capi.Handle(n, 1, 16)
    h.set_params(prm.build_params())
    sched = np.zeros((n, prm.MAXSCHED), np.int32)
    sched[:, 0] = (7 << 16) | (T[:, 3].astype(np.int32) << 8) | 55
    sched[:, 1] = (9 << 16) | (9 << 8) | 55                       # a later entry that must stay
    h.set_state(seir=T[:, 0].astype(np.int8), inf_type=T[:, 1].astype(np.int8), inf_sev=T[:, 2].astype(np.int8),
                sched=sched, sched_cnt=np.full(n, 2, np.uint8))
    h.advance_seir(7, None)
    d = h.get_state()
    assert np.array_equal(d["seir"], T[:, 4].astype(np.int8))
    assert np.array_equal(d["inf_type"], T[:, 5].astype(np.int8))
    assert np.array_equal(d["inf_sev"], T[:, 6].astype(np.int8))
    assert np.all(d["sched_cnt"] == 1) and np.all(d["sched"][:, 0] == sched[:, 1])
    h.close()


def test_epidemiology_table_helpers():
    """Epidemiology.convert_celltype_to_base_infection_prob / convert_simcelltype_to_contact_tracing_success_prob /
    update_quarantine (epidemiology.py:627-630, 978-1014) against the parameter file"""
    from types import SimpleNamespace
    from mtdcovabm_b200 import params as prm
    from mtdcovabm_b200.epidemiology import Epidemiology
    ep = prm.load_json("epidemiology.json")
    e = Epidemiology(ep, 10, 0, 1.0, None, {}, SimpleNamespace(), {}, {}, {}, SimpleNamespace())
    hh, wp, sc, cm = e.params.p_household, e.params.p_workplace, e.params.p_school, e.params.p_community
    assert e.convert_celltype_to_base_infection_prob(5, 5, celltype=CellType.Household) == hh
    assert e.convert_celltype_to_base_infection_prob(5, 5, celltype=CellType.Workplace) == wp       # lives there
    assert e.convert_celltype_to_base_infection_prob(5, 6, celltype=CellType.Workplace) == cm       # quirk: visitor
    assert e.convert_celltype_to_base_infection_prob(5, 6, celltype=CellType.Classroom) == sc
    assert e.convert_celltype_to_base_infection_prob(5, 6, celltype=CellType.Transport) == cm
    assert [e.convert_simcelltype_to_contact_tracing_success_prob(k) for k in range(4)] == list(ep["contact_tracing_success_probability"])
    agent = {"quarantine_days": None}
    e.update_quarantine(agent, 3, 50, 17, 60)
    assert agent["quarantine_days"] == [3, 50, 17]


def test_initialize_agent_states_kat():
    """seirstateutil.initialize_agent_states (seirstateutil.py:5-32) on the reference's legacy RNG streams
    (tests/golden/make_cells_kat.py -> init_states_kat.npz): the host mirror assigns exactly the same state to every id"""
    import random
    from mtdcovabm_b200 import seirstateutil as ssu
    from mtdcovabm_b200.epidemiologyclasses import SEIRState
    K = np.load(os.path.join(os.path.dirname(__file__), "golden", "init_states_kat.npz"))
    k = 0
    while "c%d_n" % k in K.files:
        n, s1, s2 = (int(x) for x in K["c%d_n" % k])
        np.random.seed(s1); random.seed(s2)
        st = ssu.initialize_agent_states(n, [float(x) for x in K["c%d_dist" % k]], {i: SEIRState.Undefined for i in range(n)})
        assert [int(st[i]) for i in range(n)] == [int(x) for x in K["c%d_out" % k]]
        k += 1
    assert k == 4


def test_schedule_helpers_match_reference_kat():
    """Epidemiology.schedule_test / schedule_quarantine / schedule_hospitalisation (epidemiology.py:502-660): the host
    mirrors return the same values, leave the same records and consume the same draws as the reference on 480 seeded
    calls each (tests/golden/make_schedule_kat.py -> schedule_kat.json).  CPU-only: these are host-side helpers."""
    import json
    import os
    from types import SimpleNamespace
    from mtdcovabm_b200 import epidemiology as E, params as prm
    from mtdcovabm_b200.epidemiologyclasses import QuarantineType
    from tests.golden import make_schedule_kat as MK

    def make(variant, dp):
        agents_epi = {}
        vu = SimpleNamespace(agents_vaccination_doses={}, agents_seir_state={}, agents_infection_type={}, agents_infection_severity={})
        return E.Epidemiology(prm.load_json(variant), 100, 0, 1.0, None, agents_epi, vu, {}, {}, {}, dp), agents_epi

    want = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "schedule_kat.json")))
    got = MK.battery(make, QuarantineType)
    assert len(got) == len(want) and sum(1 for k in want if k[0] == "test" and k[1][0]) > 50
    for g, w in zip(got, want):
        assert g == w
