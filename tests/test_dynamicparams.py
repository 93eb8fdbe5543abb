"""DynamicParams mirror (mtdcovabm_b200/dynamicparams.py) against known answers of the reference's
DynamicParams.refresh_dynamic_parameters over 60 days x 3 parameter files x {day, infectious-rate} thresholds
(tests/golden/make_dynparams_kat.py), and -- on the GPU -- fed by the device occupancy histogram (K6)."""
import json
import os

import numpy as np
import pytest

from mtdcovabm_b200.dynamicparams import DynamicParams, seir_states_from_occupancy

KAT = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "dynparams_kat.json")))
FIELDS = ["masks_hygiene_distancing_multiplier", "vaccination_propensity", "quarantine_enabled", "testing_enabled",
          "contact_tracing_enabled", "travel_restrictions_multiplier", "workplaces_lockdown", "schools_lockdown",
          "entertainment_lockdown", "activities_lockdown", "airport_lockdown"]


@pytest.mark.parametrize("key", sorted(KAT))
def test_refresh_matches_reference(key):
    dp = DynamicParams(1000, 2000, 40, KAT[key]["params"])
    for row in KAT[key]["rows"]:
        dp.refresh_dynamic_parameters(row["day"], 40, 3, 2, 1, seir_states=tuple(row["seir"]))
        assert dp.statistics.infectious_rate == row["infectious_rate"]
        assert dp.statistics.new_infectious == row["new_infectious"]
        for f in FIELDS:
            assert getattr(dp, f) == row[f], (key, row["day"], f)


def test_refresh_without_counts_fails_loudly():
    dp = DynamicParams(10, 0, 0, KAT[sorted(KAT)[0]]["params"])
    with pytest.raises(RuntimeError):
        dp.refresh_dynamic_parameters(1)


@pytest.mark.gpu
def test_refresh_from_device_occupancy():
    from mtdcovabm_b200 import capi, params as prm
    n = 5003
    rng = np.random.default_rng(2)
    seir = rng.integers(0, 6, n).astype(np.int8)
    h = capi.Handle(n, 1, 16)
    h.set_params(prm.build_params())
    h.set_state(seir=seir)
    occ = h.occupancy()
    assert [int(x) for x in occ] == [int((seir == k).sum()) for k in range(6)]
    key = "epidemiology_strict_days.json:rate"
    a = DynamicParams(n, 0, 0, KAT[key]["params"]); b = DynamicParams(n, 0, 0, KAT[key]["params"])
    a.refresh_dynamic_parameters(12, handle=h)
    b.refresh_dynamic_parameters(12, seir_states=seir_states_from_occupancy([int((seir == k).sum()) for k in range(6)]))
    assert a.statistics.infectious_rate == b.statistics.infectious_rate > 0
    assert all(getattr(a, f) == getattr(b, f) for f in FIELDS)
    h.close()
