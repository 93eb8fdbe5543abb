"""Edge cases of the day step on the GPU, checked against the oracle (which is pinned on the reference's golden days):
empty day, cells of one interval / one agent, inclusive touching intervals, zero overlap, every size-class boundary up
to the largest shared-memory cell (1024 intervals) and beyond (global-workspace class, up to 16384 intervals), shuffled arrival order, indid == 9 and Classroom cells, and the loud
failures (oversized cell, row capacity, invalid cell id)."""
import numpy as np
import pytest

from mtdcovabm_b200 import capi
from mtdcovabm_b200 import params as prm
from oracle import mtd_oracle as mo
from tests import helpers as H

pytestmark = pytest.mark.gpu
P = prm.build_params()


def population(n_agents, n_cells, seed=0, cell_type=None, cell_indid=None):
    rng = np.random.default_rng(seed)
    st = mo.AgentState(n_agents)
    st.age_bracket[:] = rng.integers(0, 16, n_agents); st.epi_age_bracket[:] = rng.integers(0, 10, n_agents)
    st.soc_rate[:] = 0.8 + 0.4 * rng.random(n_agents); st.res_cellid[:] = rng.integers(0, n_cells, n_agents)
    st.seir[:] = np.where(rng.random(n_agents) < 0.3, 3, 1)
    inf = st.seir == 3
    st.inf_type[inf] = 3; st.inf_sev[inf] = 1
    ct = np.ascontiguousarray(cell_type if cell_type is not None else (np.arange(n_cells) % 11), np.int8)
    ci = np.ascontiguousarray(cell_indid if cell_indid is not None else np.full(n_cells, -1), np.int16)
    return st, ct, ci


def run_both(st, ct, ci, iv, day=9, seed=5, **kw):
    cell, agent, start, end = [np.ascontiguousarray(x, t) for x, t in zip(iv, (np.int32, np.int32, np.uint8, np.uint8))]
    h = H.make_handle(P, st, ct, ci, max(len(cell), 1) + 16, **kw)
    d = h.make_day(capi.MODE_PHILOX, day, 0.1, True, True, seed=seed)
    h.simulate_contact_network(d, agent, cell, start, end)
    h.simulate_direct_contacts(d)
    c = h.counters()
    grow, gdur, gcell = h.get_contacts()
    gexp, gst = h.get_exposed(), h.get_state()
    h.close()
    ost = st.copy()
    ctx = mo.make_ctx(mo.MODE_PHILOX, day, 0.1, True, True, seed=seed)
    orow, odur, orc, s1 = mo.contact_network(P, ctx, cell, agent, start, end, ct, ci, ost)
    oexp, _, _ = mo.transmission(P, ctx, cell, agent, start, end, ct, orow, odur, orc, ost)
    assert np.array_equal(grow, orow) and np.array_equal(gdur, odur) and np.array_equal(gcell, orc)
    assert c.pairs_evaluated == s1.pairs_evaluated and c.n_potential == s1.n_potential
    assert np.array_equal(gexp, oexp) and H.state_diff(gst, ost) == []
    return grow, gdur, c


def test_empty_day():
    st, ct, ci = population(10, 4)
    h = H.make_handle(P, st, ct, ci, 16)
    d = h.make_day(capi.MODE_PHILOX, 1, 0.0, seed=1)
    z = np.zeros(0, np.int32)
    h.simulate_contact_network(d, z, z, np.zeros(0, np.uint8), np.zeros(0, np.uint8))
    h.simulate_direct_contacts(d)
    c = h.counters()
    assert c.error == 0 and c.n_rows == 0 and c.pairs_evaluated == 0 and len(h.get_contacts()[0]) == 0
    assert len(h.get_exposed()) == 0 and H.state_diff(h.get_state(), st) == []
    h.close()


def test_hand_made_cells():
    """one interval; one agent twice; touching (end == start, inclusive -> 1 timestep); disjoint; repeated agent + partner"""
    st, ct, ci = population(12, 8, cell_type=[0, 0, 0, 0, 1, 6, 0, 0], cell_indid=[-1, -1, -1, -1, 9, -1, -1, -1])
    iv = ([0, 1, 1, 2, 2, 3, 3, 4, 4, 4, 5, 5, 5, 6, 6, 6],           # cell
          [0, 1, 1, 2, 3, 4, 5, 6, 7, 6, 8, 9, 10, 11, 0, 11],        # agent
          [0, 0, 50, 10, 40, 10, 41, 0, 20, 60, 5, 5, 5, 0, 30, 90],  # start
          [143, 20, 70, 40, 80, 40, 80, 30, 70, 100, 100, 100, 100, 40, 100, 120])
    rows, dur, c = run_both(st, ct, ci, iv)
    by_cell = {}
    for (s, a, b, s0, e0), du in zip(rows.tolist(), dur.tolist()):
        by_cell.setdefault((a, b), (s0, e0, du))
    assert (2, 3) in by_cell and by_cell[(2, 3)] == (40, 40, 1)                  # touching intervals: one timestep
    assert not any(4 in k and 5 in k for k in by_cell)                          # [10,40] and [41,80] never meet
    assert not any(k[0] == k[1] for k in by_cell)                               # an agent is never its own contact
    if (11, 0) in by_cell or (0, 11) in by_cell:                                # repeated agent: first start / last end
        s0, e0, du = by_cell.get((11, 0), by_cell.get((0, 11)))
        assert (s0, e0, du) == (30, 100, 11 + 11)


@pytest.mark.parametrize("m", [1, 2, 8, 9, 16, 17, 32, 33, 50, 64, 65, 128, 129, 192, 193, 320, 321, 700, 1024, 1025, 2048, 5000,
                               16384])
def test_size_class_boundaries(m):
    """one cell of exactly m intervals (+ a few small neighbours), arrival order shuffled by the scatter"""
    rng = np.random.default_rng(m)
    n_agents = m + 40
    st, ct, ci = population(n_agents, 6, seed=m, cell_type=[1, 0, 5, 8, 4, 0])
    agents = rng.permutation(n_agents)[:m]
    if m > 4:
        agents[rng.integers(0, m, m // 8 if m <= 1024 else m // 64)] = agents[0]  # some agents come back several times
        agents[rng.integers(0, m, m // 8)] = agents[rng.integers(0, m, m // 8)]  # ... and others once or twice more
    start = rng.integers(0, 120, m); end = np.minimum(143, start + rng.integers(0, 60, m))
    cell = np.zeros(m, np.int64)
    extra_c = rng.integers(1, 6, 30); extra_a = rng.integers(0, n_agents, 30)
    extra_s = rng.integers(0, 100, 30); extra_e = extra_s + rng.integers(0, 43, 30)
    o = rng.permutation(m + 30)                                                  # interleave the streams
    iv = [np.concatenate(x)[o] for x in ((cell, extra_c), (agents, extra_a), (start, extra_s), (end, extra_e))]
    rows, dur, c = run_both(st, ct, ci, iv, seed=m)
    assert c.error == 0
    if m >= 8:
        assert len(rows) > 0


def test_oversized_cell_fails_loudly():
    m = capi.MAX_CELL_INTERVALS + 1                # the reference has no limit; this library's is 16384 intervals per cell
    st, ct, ci = population(m, 2)
    h = H.make_handle(P, st, ct, ci, m + 16)
    d = h.make_day(capi.MODE_PHILOX, 1, 0.0, seed=1)
    h.simulate_contact_network(d, np.arange(m, dtype=np.int32), np.zeros(m, np.int32), np.zeros(m, np.uint8),
                               np.full(m, 100, np.uint8))
    with pytest.raises(RuntimeError):
        h.counters()
    h.close()


def test_row_capacity_fails_loudly():
    st, ct, ci = population(200, 2, cell_type=[1, 1])
    h = H.make_handle(P, st, ct, ci, 216, max_rows=8)
    d = h.make_day(capi.MODE_PHILOX, 1, 0.0, seed=1)
    h.simulate_contact_network(d, np.arange(200, dtype=np.int32), np.zeros(200, np.int32), np.zeros(200, np.uint8),
                               np.full(200, 100, np.uint8))
    with pytest.raises(RuntimeError):
        h.counters()
    h.close()


def test_invalid_cell_id_fails_loudly():
    st, ct, ci = population(4, 2)
    h = H.make_handle(P, st, ct, ci, 16)
    d = h.make_day(capi.MODE_PHILOX, 1, 0.0, seed=1)
    h.simulate_contact_network(d, np.arange(4, dtype=np.int32), np.array([0, 1, 7, 0], np.int32), np.zeros(4, np.uint8),
                               np.full(4, 9, np.uint8))
    with pytest.raises(RuntimeError):
        h.counters()
    h.close()


@pytest.mark.parametrize("n_rep_iv", [5, 24])
def test_exposure_timestep_with_many_intervals_per_agent(n_rep_iv):
    """an infectious agent that comes back to the same cell n times (5: staged in shared memory; 24: more than the
    staging holds -> record walk) beside susceptible partners with several visits each, arrival order shuffled: the
    exposure timestep is drawn from the flattened overlap list in LIST order, so every schedule entry must match"""
    rng = np.random.default_rng(n_rep_iv)
    n_cells, per = 400, 6
    n_agents = n_cells * per
    st, ct, ci = population(n_agents, n_cells, seed=3, cell_type=np.zeros(n_cells))
    st.seir[:] = 1; st.inf_type[:] = 0; st.inf_sev[:] = 0
    cell, agent, start, end = [], [], [], []
    for c_ in range(n_cells):
        a0 = c_ * per
        st.seir[a0] = 3; st.inf_type[a0] = 4; st.inf_sev[a0] = 1; st.res_cellid[a0:a0 + per] = c_
        s0 = np.sort(rng.choice(140, n_rep_iv, replace=False))
        for s in rng.permutation(s0):                                       # the reference keeps LIST order, not time order
            cell.append(c_); agent.append(a0); start.append(s); end.append(min(143, s + rng.integers(0, 4)))
        for k in range(1, per):
            for _ in range(3):
                s = int(rng.integers(0, 100))
                cell.append(c_); agent.append(a0 + k); start.append(s); end.append(s + int(rng.integers(5, 43)))
    o = rng.permutation(len(cell))
    o = o[np.argsort(np.array(cell)[o] % 7, kind="stable")]               # cells interleaved, per-cell order scrambled
    iv = [np.array(x)[o] for x in (cell, agent, start, end)]
    rows, dur, c = run_both(st, ct, ci, iv, seed=n_rep_iv)
    assert c.error == 0 and c.n_exposed > 20
