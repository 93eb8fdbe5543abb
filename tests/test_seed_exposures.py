"""K9 -- Epidemiology.simulate_seir_state_transition called for exposures BEFORE day 1 (the itinerary's initial cases,
itinerary.py:373-395; epidemiology.py:233-500 with reset_states): 1200 calls of the unmodified reference with recorded
draws (tests/golden/seed_kat.npz, made by tests/golden/make_seed_kat.py) replayed through the C oracle (CPU) and the CUDA
path (GPU), plus Philox-mode equality CUDA == oracle."""
import os

import numpy as np
import pytest

from mtdcovabm_b200 import params as prm
from oracle import mtd_oracle as mo

HERE = os.path.dirname(os.path.abspath(__file__))
NONE = prm.NONE


def _kat():
    return np.load(os.path.join(HERE, "golden", "seed_kat.npz"))


def _initial_state(z):
    n = len(z["cur_day"])
    st = mo.AgentState(n)
    st.epi_age_bracket[:] = z["bracket"]
    st.seir[:] = 3                                   # the callers' agents start Infectious
    st.test_day[:] = z["in_test"]; st.test_result_day[:] = z["in_tres"]; st.hosp[:] = z["in_hosp"]
    return st


def _groups(z):
    """cases that share one day context (current day, testing / quarantine switches) form one batch call"""
    key = z["cur_day"].astype(np.int64) * 4 + z["testing"] * 2 + z["quar"]
    for k in np.unique(key):
        ids = np.nonzero(key == k)[0].astype(np.int32)
        yield int(z["cur_day"][ids[0]]), bool(z["testing"][ids[0]]), bool(z["quar"][ids[0]]), ids


def _check_against_reference(z, st, out_by_id):
    n = len(z["cur_day"])
    for i in range(n):
        o = out_by_id[i]
        assert tuple(o[:4]) == tuple(z["out_state"][i]), "case %d: states / recovered" % i
        assert (st.seir[i], st.inf_type[i], st.inf_sev[i]) == tuple(z["out_state"][i][:3])
        ns = int(z["out_nsched"][i])
        assert st.sched_cnt[i] == ns == o[4], "case %d: schedule length" % i
        got = [mo.unpack_sched(v) for v in st.sched[i, :ns]]
        assert got == [tuple(int(x) for x in e) for e in z["out_sched"][i, :ns]], "case %d: schedule entries" % i
    assert np.array_equal(st.test_day, z["out_test"])
    assert np.array_equal(st.test_result_day, z["out_tres"])
    assert np.array_equal(st.hosp, z["out_hosp"])
    assert (st.quarantine == NONE).all()             # schedule_quarantine's record never survives the caller (:390)


def test_oracle_replays_the_reference_calls():
    z = _kat()
    P = prm.build_params()
    st = _initial_state(z)
    out_by_id = {}
    for day, te, qe, ids in _groups(z):
        ctx = mo.make_ctx(mo.MODE_REPLAY, day, 0.0, te, qe)
        out, stats = mo.seed_exposures(P, ctx, st, ids, exposure_day=z["expo_day"][ids], asym_mult=z["asym"][ids],
                                       draws=z["draws"][ids])
        assert stats.tape_miss == 0 and stats.sched_overflow == 0
        out_by_id.update({int(a): out[k] for k, a in enumerate(ids)})
    _check_against_reference(z, st, out_by_id)


def test_oracle_philox_exposure_day_range_and_determinism():
    P = prm.build_params()
    n = 4000
    st = mo.AgentState(n); st.epi_age_bracket[:] = np.arange(n) % 10; st.seir[:] = 3
    ids = np.arange(n, dtype=np.int32)
    ctx = mo.make_ctx(mo.MODE_PHILOX, 1, 0.0, False, False, seed=9)
    out, _ = mo.seed_exposures(P, ctx, st, ids, pre_exposure_days=7)
    assert out[:, 6].min() == -6 and out[:, 6].max() == 0          # arange(simday - 7, simday), itinerary.py:375
    assert 0 <= out[:, 5].min() and out[:, 5].max() <= 143
    st2 = mo.AgentState(n); st2.epi_age_bracket[:] = st.epi_age_bracket; st2.seir[:] = 3
    out2, _ = mo.seed_exposures(P, mo.make_ctx(mo.MODE_PHILOX, 1, 0.0, False, False, seed=9), st2, ids[::-1].copy(), pre_exposure_days=7)
    assert np.array_equal(out2[::-1], out) and all(st.dyn_equal(st2).values())    # keyed by agent, not by position
    # every seeded agent is in a state the reset table allows, and nobody is left Susceptible / Undefined
    assert set(np.unique(st.seir)) <= {2, 3, 4}


@pytest.mark.gpu
def test_gpu_replays_the_reference_calls():
    from tests import helpers as H
    z = _kat()
    P = prm.build_params()
    st = _initial_state(z)
    h = H.make_handle(P, st, np.zeros(1, np.int8), np.full(1, -1, np.int16), 16)
    out_by_id = {}
    for day, te, qe, ids in _groups(z):
        from mtdcovabm_b200 import capi
        d = h.make_day(capi.MODE_REPLAY, day, 0.0, te, qe)
        out = h.seed_exposures(d, ids, exposure_day=z["expo_day"][ids], asym_mult=z["asym"][ids], draws=z["draws"][ids])
        out_by_id.update({int(a): out[k] for k, a in enumerate(ids)})
    dev = h.get_state()
    got = mo.AgentState(st.n)
    for f in ("seir", "inf_type", "inf_sev", "sched", "sched_cnt", "test_day", "test_result_day", "quarantine", "hosp"):
        getattr(got, f)[...] = dev[f]
    _check_against_reference(z, got, out_by_id)
    h.close()


@pytest.mark.gpu
@pytest.mark.parametrize("day,te,qe,seed", [(1, False, False, 3), (1, True, True, 4), (4, True, False, 5)])
def test_gpu_philox_equals_oracle(day, te, qe, seed):
    from mtdcovabm_b200 import capi
    from tests import helpers as H
    P = prm.build_params()
    n = 20000
    rs = np.random.RandomState(seed)
    st = mo.AgentState(n)
    st.epi_age_bracket[:] = rs.randint(0, 10, n); st.seir[:] = 1
    ids = np.sort(rs.choice(n, 6000, replace=False)).astype(np.int32)
    st.seir[ids] = 3
    has_test = ids[::9]
    st.test_day[has_test] = (day + 2, 60); st.test_result_day[has_test] = (day + 3, 60)
    st.hosp[ids[::11]] = (day - 1, 50, day + 6)
    asym = np.where(rs.rand(len(ids)) < 0.2, 0.1, 1.0)
    h = H.make_handle(P, st, np.zeros(1, np.int8), np.full(1, -1, np.int16), 16)
    d = h.make_day(capi.MODE_PHILOX, day, 0.0, te, qe, seed=seed)
    out = h.seed_exposures(d, ids, pre_exposure_days=14, asym_mult=asym)
    ost = st.copy()
    oout, _ = mo.seed_exposures(P, mo.make_ctx(mo.MODE_PHILOX, day, 0.0, te, qe, seed=seed), ost, ids, pre_exposure_days=14,
                                asym_mult=asym)
    assert np.array_equal(out, oout)
    assert H.state_diff(h.get_state(), ost) == []
    # explicit exposure days + a caller-supplied timestep list (the per-call form of the reference's signature)
    h2 = H.make_handle(P, st, np.zeros(1, np.int8), np.full(1, -1, np.int16), 16)
    ed = (day - rs.randint(0, 30, len(ids))).astype(np.int32)
    tsl = np.arange(42, 121, dtype=np.uint8)
    out2 = h2.seed_exposures(d, ids, exposure_day=ed, timesteps=tsl)
    ost2 = st.copy()
    oout2, _ = mo.seed_exposures(P, mo.make_ctx(mo.MODE_PHILOX, day, 0.0, te, qe, seed=seed), ost2, ids, exposure_day=ed, timesteps=tsl)
    assert np.array_equal(out2, oout2) and H.state_diff(h2.get_state(), ost2) == []
    assert out2[:, 5].min() >= 42 and out2[:, 5].max() <= 120
    h.close(); h2.close()


@pytest.mark.gpu
def test_gpu_rejects_bad_ids_and_missing_draws():
    from mtdcovabm_b200 import capi
    from tests import helpers as H
    P = prm.build_params()
    st = mo.AgentState(100)
    h = H.make_handle(P, st, np.zeros(1, np.int8), np.full(1, -1, np.int16), 16)
    d = h.make_day(capi.MODE_PHILOX, 1, 0.0, False, False)
    with pytest.raises(RuntimeError):
        h.seed_exposures(d, np.array([5, 100], np.int32), pre_exposure_days=7)
    with pytest.raises(RuntimeError):
        h.seed_exposures(d, np.array([5], np.int32))                                   # no exposure day, no range
    with pytest.raises(RuntimeError):
        h.seed_exposures(h.make_day(capi.MODE_REPLAY, 1, 0.0, False, False), np.array([5], np.int32), exposure_day=[0])
    h.close()


@pytest.mark.gpu
def test_dropin_simulate_seir_state_transition_matches_reference_calls():
    """the reference's own signature and return tuple, one call per case (replay draws), on a spread of the KAT cases"""
    from types import SimpleNamespace
    from mtdcovabm_b200 import bridge
    from mtdcovabm_b200.epidemiology import Epidemiology
    from mtdcovabm_b200.epidemiologyclasses import InfectionType, SEIRState, Severity
    z = _kat()
    ep = prm.load_json("epidemiology.json")
    dp = SimpleNamespace(quarantine_enabled=False, testing_enabled=False, contact_tracing_enabled=False,
                         masks_hygiene_distancing_multiplier=0.0)
    n = len(z["cur_day"])
    e = Epidemiology(ep, n, 0, 1.0, None, {}, SimpleNamespace(), {}, {}, {}, dp, mode="replay")
    lst = lambda a: None if a[0] == NONE else [int(x) for x in a]
    picked = list(range(0, n, 23)) + [int(i) for i in np.nonzero(z["out_test"][:, 0] != z["in_test"][:, 0])[0][:12]]
    for i in picked:
        dp.testing_enabled, dp.quarantine_enabled = bool(z["testing"][i]), bool(z["quar"][i])
        agent = {"state_transition_by_day": None, "test_day": lst(z["in_test"][i]), "test_result_day": lst(z["in_tres"][i]),
                 "hospitalisation_days": lst(z["in_hosp"][i]), "quarantine_days": None}
        e.tape = z["draws"][i]
        trans, seir, it, sv, rec = e.simulate_seir_state_transition(int(z["cur_day"][i]), agent, i, int(z["expo_day"][i]),
                                                                    np.arange(144), [], int(z["bracket"][i]), None,
                                                                    float(z["asym"][i]))
        assert (seir, it, sv, int(rec)) == tuple(z["out_state"][i]) and isinstance(seir, SEIRState)
        assert isinstance(it, InfectionType) and isinstance(sv, Severity)
        ns = int(z["out_nsched"][i])
        assert [[int(x) for x in t] for t in trans] == [[int(x) for x in t] for t in z["out_sched"][i, :ns]]
        assert agent["test_day"] == lst(z["out_test"][i]) and agent["test_result_day"] == lst(z["out_tres"][i])
        assert agent["hospitalisation_days"] == lst(z["out_hosp"][i]) and agent["quarantine_days"] is None
    bridge.close_handles()


@pytest.mark.gpu
def test_dropin_seed_initial_cases_batch():
    """itinerary.py:373-395 as one batch: every agent that starts Infectious leaves with a schedule (or a finished
    trajectory), a state the reset table allows, and its quarantine record untouched"""
    from types import SimpleNamespace
    from mtdcovabm_b200 import bridge
    from mtdcovabm_b200.epidemiology import Epidemiology
    n = 3000
    rs = np.random.RandomState(5)

    class Static:
        def get(self, a, name):
            return {"age_bracket_index": a % 16, "epi_age_bracket_index": a % 10, "soc_rate": 1.0, "res_cellid": 0}[name]

    agents_epi = {a: {"state_transition_by_day": None, "test_day": None, "test_result_day": None, "hospitalisation_days": None,
                      "quarantine_days": [2, 50, 12] if a % 50 == 0 else None} for a in range(n)}
    vu = SimpleNamespace(agents_seir_state={a: 1 for a in range(n)}, agents_infection_type={}, agents_infection_severity={},
                         agents_vaccination_doses={})
    ids = [int(a) for a in np.sort(rs.choice(n, 400, replace=False))]
    for a in ids:
        vu.agents_seir_state[a] = 3
    dp = SimpleNamespace(quarantine_enabled=True, testing_enabled=True, contact_tracing_enabled=False,
                         masks_hygiene_distancing_multiplier=0.0)
    e = Epidemiology(prm.load_json("epidemiology.json"), n, 0, 1.0, Static(), agents_epi, vu, {}, {}, {}, dp, seed=11)
    assert e.seed_initial_cases(1, ids, 7) == ids
    seen = set()
    for a in ids:
        s, it, sv = int(vu.agents_seir_state[a]), int(vu.agents_infection_type[a]), int(vu.agents_infection_severity[a])
        seen.add((s, it, sv))
        assert (s, it, sv) in {(2, 1, 0), (3, 2, 0), (3, 3, 0), (3, 4, 1), (3, 4, 2), (3, 4, 3), (4, 0, 0), (2, 2, 0)}
        ent = agents_epi[a]["state_transition_by_day"]
        assert all(d >= 1 for d, _, _ in ent) and [d for d, _, _ in ent] == sorted(d for d, _, _ in ent)
        assert agents_epi[a]["quarantine_days"] == ([2, 50, 12] if a % 50 == 0 else None)
    assert len(seen) >= 4 and all(vu.agents_seir_state[a] == 1 for a in range(n) if a not in set(ids))
    bridge.close_handles()
