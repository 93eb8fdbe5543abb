"""The tensor-native day loop with every stage of the state step on the device (mtdcovabm_b200/dayloop.py): 40 days of a
20k-agent tiled population under the strict-interventions parameter file (BASELINE.json config 5 shape: testing,
quarantine, contact tracing, masks, vaccination).  Invariants: the population is conserved, nothing overflows, every
stage does work, and the run is reproducible bit for bit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(seed):
    import torch
    from mtdcovabm_b200 import dayloop, params as prm, synth
    np.random.seed(seed)
    bank = synth.load_bank()
    pop = synth.SynthPopulation(bank, 20, torch.device("cuda", 0), seed=seed, infectious_frac=0.01)
    ep = dict(prm.load_json("epidemiology_strict_days.json"))
    ep["vaccination_day_thresholds"] = [[0, 0.0], [10, 0.4]]           # vaccination starts on day 10 in this test
    ep["vaccination_infectiousrate_thresholds"] = []
    run = dayloop.EpidemicRun(pop, ep, seed=seed)
    log = [run.day(d) for d in range(1, 41)]
    st = run.h.get_state()
    last, rec = run.h.get_vaccination()
    n = run.n_present
    run.close()
    return log, st, last, n


def test_full_state_step_day_loop():
    log, st, last, n = _run(3)
    for r in log:
        assert sum(r["occupancy"][1:]) == n and r["sched_overflow"] == 0
    assert sum(r["exposed"] for r in log) > 100
    assert sum(r["traced"] for r in log) > 0 and sum(r["primary"] for r in log) > 0
    assert sum(r["vaccinations"] for r in log) > 0 and all(r["vaccinations"] == 0 for r in log[:9])
    assert log[5]["testing"] and log[5]["quarantine"] and log[5]["tracing"] and log[5]["mhd"] > 0
    from mtdcovabm_b200 import params as prm
    dosed = (last != prm.NONE) & (st["seir"] != 0)
    assert 0 < dosed.sum() <= sum(r["vaccinations"] for r in log)
    assert (st["quarantine"][:, 0] != prm.NONE).sum() > 0 and (st["test_day"][:, 0] != prm.NONE).sum() > 0
    # records expire as in the itinerary stage (itinerary.py:2396, 2456, 2486): nothing stale is left behind, so agents
    # can be tested and quarantined again later in the run
    q, hp, td, tr = st["quarantine"], st["hosp"], st["test_day"], st["test_result_day"]
    assert ((hp[:, 0] != prm.NONE) & (hp[:, 2] < 40) & (hp[:, 0] != hp[:, 2])).sum() == 0
    assert ((q[:, 0] != prm.NONE) & (q[:, 2] < 40) & (q[:, 0] != q[:, 2])).sum() <= 5      # (hospital discharge on the end day keeps it)
    assert ((td[:, 0] != prm.NONE) & (td[:, 0] < 40 - 6)).sum() == 0
    assert ((tr[:, 0] != prm.NONE) & (tr[:, 0] < 40)).sum() == 0
    log2, st2, last2, _ = _run(3)
    assert [r["occupancy"] for r in log] == [r["occupancy"] for r in log2]
    assert [r["primary"] for r in log] == [r["primary"] for r in log2]
    assert all(np.array_equal(st[k], st2[k]) for k in st) and np.array_equal(last, last2)
