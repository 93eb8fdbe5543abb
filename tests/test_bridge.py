"""Host-side marshalling of the drop-in classes (mtdcovabm_b200/bridge.py): the vectorised flatten of the reference's
containers must equal the plain per-element restatement, with and without the static-attribute cache.  CPU only."""
import numpy as np
import pytest

from mtdcovabm_b200 import bridge
from tests.golden import golden_io as gio
from tests.test_dropin_api import FakeStatic, reference_containers

FIELDS = ("seir", "inf_type", "inf_sev", "last_dose_day", "sched", "sched_cnt", "test_day", "test_result_day", "quarantine", "hosp")
STATIC = ("age_bracket", "epi_age_bracket", "soc_rate", "res_cellid")


def _day(name, k):
    z = gio.load_run(name)
    return gio.load_day(z, int(z["days"][k]))


@pytest.mark.parametrize("name", ["ref_1k_seed6", "ref_1k_strict_seed7", "ref_10k_strict_seed12"])
def test_flatten_intervals_equals_the_dict_order_loop(name):
    g = _day(name, 1)
    _, vu = reference_containers(g, g.pre)
    c, a, s, e = bridge.flatten_intervals(vu.cells_agents_timesteps)
    assert (c.dtype, a.dtype, s.dtype, e.dtype) == (np.int32, np.int32, np.uint8, np.uint8)
    want = [(int(cell), int(x), int(y), int(w)) for cell, lst in vu.cells_agents_timesteps.items() for (x, y, w) in lst]
    assert list(zip(c.tolist(), a.tolist(), s.tolist(), e.tolist())) == want
    assert np.array_equal(c, g.iv_cell) and np.array_equal(a, g.iv_agent)
    empty = bridge.flatten_intervals({})
    assert all(len(x) == 0 for x in empty)
    ragged = bridge.flatten_intervals({7: [], 3: [(1, 2, 3)], 9: [], 4: [(5, 6, 7), (8, 9, 10)]})
    assert ragged[0].tolist() == [3, 4, 4] and ragged[1].tolist() == [1, 5, 8] and ragged[3].tolist() == [3, 7, 10]


@pytest.mark.parametrize("name", ["ref_1k_seed6", "ref_1k_strict_seed7", "ref_10k_strict_seed12"])
def test_flatten_equals_the_per_element_restatement(name):
    bridge.clear_static_cache()
    static = None
    for k in (0, 2, 3):                     # several days through ONE Static object, as the reference does: the cache fills
        g = _day(name, k)
        epi, vu = reference_containers(g, g.pre)
        static = static or FakeStatic(g.pre)
        active = np.unique(g.iv_agent)
        want = bridge.flatten_reference_loop(g.N, static, epi, vu, active)
        for n_locals in (None, g.n_locals):
            got = bridge.flatten(g.N, static, epi, vu, active, n_locals=n_locals)
            for f in FIELDS:
                assert np.array_equal(getattr(got, f), getattr(want, f)), (f, k, n_locals)
            for f in STATIC:                # the device reads the static attributes of the day's active agents only
                assert np.array_equal(getattr(got, f)[active], getattr(want, f)[active]), (f, k, n_locals)


def test_static_cache_keeps_locals_only_and_is_per_static_object():
    bridge.clear_static_cache()
    g = _day("ref_1k_seed6", 0)
    epi, vu = reference_containers(g, g.pre)
    active = np.unique(g.iv_agent)

    class Counting(FakeStatic):
        calls = 0

        def get(self, a, name):
            Counting.calls += 1
            return super().get(a, name)

    s1 = Counting(g.pre)
    bridge.flatten(g.N, s1, epi, vu, active, n_locals=g.n_locals)
    first = Counting.calls
    assert first == 4 * len(active)
    bridge.flatten(g.N, s1, epi, vu, active, n_locals=g.n_locals)
    n_tour = int((active >= g.n_locals).sum())
    assert Counting.calls - first == 4 * n_tour          # tourists are re-read every day (tourism.py rewrites them on arrival)
    s2 = Counting(g.pre)                                 # another Static object: nothing is taken from the first one's cache
    before = Counting.calls
    bridge.flatten(g.N, s2, epi, vu, active, n_locals=g.n_locals)
    assert Counting.calls - before == 4 * len(active)
