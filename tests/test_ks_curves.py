"""Philox mode vs the unmodified reference, statistically (north_star): daily incidence and SEIR occupancy over 32
seeds must be indistinguishable (two-sample KS p > 0.05 on every day).  The reference's 32 runs and the itineraries
they produced are in tests/golden/curves_1k.npz (tests/golden/make_curves.py); our 32 runs re-use those itineraries
and the arrivals / entry infections the itinerary stage injected, with Philox draws for everything on the hot path."""
import numpy as np
import pytest

from tests import curves as C

ALPHA = 0.05


def _run_all(runner, n=None, name="curves_1k"):
    z = C.load(name); P = C.params_for(name)
    seeds = [int(s) for s in z["seeds"]][:n]
    oi, oo, ri, ro = [], [], [], []
    for s in seeds:
        sd = C.SeedData(z, s)
        inc, occ = runner(sd, P, s)
        oi.append(inc); oo.append(occ); ri.append(sd.ref_incidence); ro.append(sd.ref_occupancy)
    return np.array(oi), np.array(oo), np.array(ri), np.array(ro)


def _assert_ks(oi, oo, ri, ro):
    assert ri.sum() > 1000 and oi.sum() > 1000              # a real epidemic on both sides (about 100 cases per run)
    ks = C.ks_table(oi, oo, ri, ro)
    for name, p in ks.items():
        assert p.min() > ALPHA, (name, p.round(3))


def test_oracle_philox_curves_match_reference_ks():
    _assert_ks(*_run_all(C.run_seed_oracle))


@pytest.mark.gpu
def test_gpu_philox_curves_match_reference_ks():
    oi, oo, ri, ro = _run_all(C.run_seed_gpu)
    _assert_ks(oi, oo, ri, ro)
    # and the CUDA path reproduces the oracle's Philox curves exactly (same draws, same arithmetic)
    ci, co, _, _ = _run_all(C.run_seed_oracle, n=4)
    assert np.array_equal(oi[:4], ci) and np.array_equal(oo[:4], co)


# ---- BASELINE.json configs 2 and 5 at their own size and parameter file: the 1k population tiled x10 so that the
# UNMODIFIED reference can run it (tests/golden/make_pop10k.py), 8 seeds x 30 days each (tests/golden/make_curves.py)
@pytest.mark.parametrize("name", ["curves_10k_normal", "curves_10k_strict"])
def test_oracle_philox_curves_match_reference_ks_10k(name):
    oi, oo, ri, ro = _run_all(C.run_seed_oracle, name=name)
    assert ri.sum() > 5000 and oi.sum() > 5000
    ks = C.ks_table(oi, oo, ri, ro)
    for k, p in ks.items():
        assert p.min() > ALPHA, (name, k, p.round(3))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["curves_10k_normal", "curves_10k_strict"])
def test_gpu_philox_curves_match_reference_ks_10k(name):
    oi, oo, ri, ro = _run_all(C.run_seed_gpu, name=name)
    ks = C.ks_table(oi, oo, ri, ro)
    for k, p in ks.items():
        assert p.min() > ALPHA, (name, k, p.round(3))
    ci, co, _, _ = _run_all(C.run_seed_oracle, n=2, name=name)
    assert np.array_equal(oi[:2], ci) and np.array_equal(oo[:2], co)
