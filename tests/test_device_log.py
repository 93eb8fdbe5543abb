"""The two per-cell logarithms of the degree formula (contactnetwork.py:237-252: math.log(x, base) = log(x) / log(avg)) are
computed with glibc's log by the reference and with CUDA's fp64 log (documented <= 1 ulp) by K1; every other logarithm on
the path has an integer argument and comes from a table the host fills with glibc's values.  A degree could only differ if
(a) the two logs differ and (b) the float degree sits within that perturbation of a rounding boundary of Python's round().
(a) is measured on the GPU over EVERY reachable argument avg = s / n (n distinct agents <= 1024, s <= 144 n total
timesteps, resp. s <= n (n - 1) potential contacts); (b) is excluded for every recorded cell: the oracle's contact rows of
all golden days and of the hand-made cells do not change when both logarithms are moved by +-1 and +-2 ulps."""
import numpy as np
import pytest

from mtdcovabm_b200 import params as prm
from oracle import mtd_oracle as mo
from tests.golden import golden_io as gio


def _rows(name, day):
    z = gio.load_run(name)
    P = prm.build_params(None, prm.load_json(gio.RUNS[name]))
    g = gio.load_day(z, day)
    ctx = mo.make_ctx(mo.MODE_REPLAY, g.day, g.mhd, g.testing_enabled, g.quarantine_enabled, tape=g.tape)
    rows, dur, rc, s1 = mo.contact_network(P, ctx, g.iv_cell, g.iv_agent, g.iv_start, g.iv_end, g.cell_type, g.cell_indid,
                                           g.pre.copy())
    return rows, s1.tape_miss


@pytest.mark.parametrize("name,day", list(gio.all_days()))
def test_no_recorded_degree_sits_within_a_log_ulp_of_a_rounding_boundary(name, day):
    try:
        base, miss = _rows(name, day)
        assert miss == 0
        for k in (-2, -1, 1, 2):
            mo.set_log_ulp_shift(k)
            rows, miss = _rows(name, day)
            assert miss == 0 and np.array_equal(rows, base), (name, day, k)
    finally:
        mo.set_log_ulp_shift(0)


@pytest.mark.gpu
def test_device_log_against_glibc_over_every_reachable_average():
    from mtdcovabm_b200 import capi
    h = capi.Handle(1, 1, 16)
    worst, n_diff, n_all = 0, 0, 0
    for n0 in range(2, 1025, 64):
        n = np.arange(n0, min(n0 + 64, 1025), dtype=np.float64)
        s = np.arange(1, 144 * int(n[-1]) + 1, dtype=np.float64)
        x = (s[None, :] / n[:, None])
        x = np.unique(x[(s[None, :] <= 144 * n[:, None]) & (x > 1.0)])           # log(avg) is only used when avg > 1
        dev = h.debug_log(x)
        ref = mo.glibc_log(x)
        ulp = np.abs(dev.view(np.int64) - ref.view(np.int64))
        worst = max(worst, int(ulp.max())); n_diff += int((ulp != 0).sum()); n_all += len(x)
    h.close()
    print("device log vs glibc: %d of %d reachable averages differ, worst %d ulp" % (n_diff, n_all, worst))
    assert worst <= 1                     # within the perturbation the +-2 ulp test above covers
