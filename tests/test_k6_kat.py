"""Pins K6 -- the record bookkeeping of Itinerary.sample_intervention_activities (itinerary.py:2283-2528): test results
due today -> contact_tracing_agent_ids + Positive quarantine, expiry of test_result_day / hospitalisation_days /
quarantine_days / test_day -- on every call the UNMODIFIED reference made in two 45-day runs with strict interventions
(tests/golden/make_k6_kat.py -> k6_kat.npz; draws of random.random() replayed).  Oracle on CPU, CUDA path on the GPU."""
import os

import numpy as np
import pytest

from mtdcovabm_b200 import params as prm
from oracle import mtd_oracle as mo

Z = np.load(os.path.join(os.path.dirname(__file__), "golden", "k6_kat.npz"))
P = prm.build_params(None, prm.load_json("epidemiology_strict_days.json"))


def groups():
    """calls sharing (day, quarantine_enabled) become the pseudo-agents 0..k-1 of one batch"""
    da, fl = Z["day_agent"], Z["flags"]
    key = da[:, 0].astype(np.int64) * 2 + fl[:, 1]
    for k in np.unique(key):
        idx = np.nonzero(key == k)[0]
        yield int(k // 2), bool(k % 2), idx


def batch_state(idx):
    st = mo.AgentState(len(idx))
    st.seir[:] = Z["seir"][idx]
    st.test_result_day[:] = Z["pre_tres"][idx]; st.test_day[:] = Z["pre_test"][idx]
    st.quarantine[:] = Z["pre_quar"][idx]; st.hosp[:] = Z["pre_hosp"][idx]
    return st


def check(idx, st_after, positive):
    for f, name in (("test_result_day", "tres"), ("test_day", "test"), ("quarantine", "quar"), ("hosp", "hosp")):
        got, exp = getattr(st_after, f) if not isinstance(st_after, dict) else st_after[f], Z["post_" + name][idx]
        assert np.array_equal(got, exp), (name, np.nonzero((got != exp).any(1))[0][:5])
    fl = Z["flags"][idx]
    fresh = fl[:, 3] == 0                                    # not yet in contact_tracing_agent_ids before the call
    pos = np.zeros(len(idx), np.int8); pos[positive] = 1
    assert np.array_equal(pos[fresh], fl[fresh, 2])
    assert np.all(pos[~fresh] >= fl[~fresh, 2])


def test_oracle_k6_matches_every_recorded_reference_call():
    n_due = n_pos = n_cleared = 0
    for day, qen, idx in groups():
        st = batch_state(idx)
        ctx = mo.make_ctx(mo.MODE_REPLAY, day, 0.0, True, qen)
        got = mo.process_test_results(P, ctx, st, None, replay_agents=np.arange(len(idx)), replay_u=Z["draws"][idx])
        assert mo.process_test_results.tape_miss == 0
        check(idx, st, got)
        n_due += int((Z["pre_tres"][idx][:, 0] == day).sum()); n_pos += len(got)
        n_cleared += int(((Z["pre_quar"][idx][:, 0] != mo.NONE) & (Z["post_quar"][idx][:, 0] == mo.NONE)).sum())
    assert n_due > 200 and n_pos > 40 and n_cleared > 20        # the fixture exercises every branch


@pytest.mark.gpu
def test_gpu_k6_matches_every_recorded_reference_call():
    from mtdcovabm_b200 import capi
    from tests import helpers as H
    for day, qen, idx in groups():
        st = batch_state(idx)
        h = capi.Handle(len(idx), 1, 16)
        h.set_params(P)
        H.upload_state(h, st)
        d = h.make_day(capi.MODE_REPLAY, day, 0.0, True, qen)
        h.set_k6_tape(np.arange(len(idx), dtype=np.int32), Z["draws"][idx])
        got = h.process_test_results(d, None)
        check(idx, h.get_state(), got)
        h.close()
