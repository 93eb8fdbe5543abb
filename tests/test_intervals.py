"""K8 interval builder (itinerary.py:1403-1488): the oracle against 400 recorded calls of the reference's
update_cell_agents_timesteps (tests/golden/make_intervals_kat.py), the CUDA builder against the oracle and the
reference, and the built stream fed straight into the day step (MTD_MEM_BUILT)."""
import os

import numpy as np
import pytest

from oracle import mtd_oracle as mo

Z = np.load(os.path.join(os.path.dirname(__file__), "golden", "intervals_kat.npz"))
ARGS = [Z[k] for k in ("ent_off", "ent_ts", "ent_cell", "grp_off", "grp_agent", "grp_rescell")]


def as_dict(a, c, s, e):
    """flat emission stream -> the reference's dict-of-lists (cells in order of first appearance)"""
    d = {}
    for ag, cell, st, en in zip(a.tolist(), c.tolist(), s.tolist(), e.tolist()):
        d.setdefault(cell, []).append((ag, st, en))
    return d


def reference_dict():
    off, rows = Z["cell_off"], Z["cell_rows"]
    return {int(c): [tuple(int(x) for x in r) for r in rows[off[i]:off[i + 1]]] for i, c in enumerate(Z["cell_order"])}


def test_oracle_builder_matches_reference_calls():
    got = as_dict(*mo.build_intervals(*ARGS, cap=100000))
    ref = reference_dict()
    assert list(got.keys()) == list(ref.keys())          # dict order of the cells = order of first appearance
    assert got == ref


@pytest.mark.gpu
def test_cuda_builder_matches_reference_and_feeds_the_day_step():
    from mtdcovabm_b200 import capi, params as prm
    n_agents, n_cells = int(Z["grp_agent"].max()) + 1, 200
    h = capi.Handle(n_agents, n_cells, 100000)
    P = prm.build_params()
    h.set_params(P)
    n = h.build_intervals(*ARGS)
    a, c, s, e = h.get_intervals(n)
    oa, oc, os_, oe = mo.build_intervals(*ARGS, cap=100000)
    assert n == len(oa) and np.array_equal(a, oa) and np.array_equal(c, oc) and np.array_equal(s, os_) and np.array_equal(e, oe)
    assert as_dict(a, c, s, e) == reference_dict()
    # the built stream runs the day without leaving the device: same rows as the stream handed in from the host
    rng = np.random.default_rng(1)
    st = mo.AgentState(n_agents)
    st.age_bracket[:] = rng.integers(0, 16, n_agents); st.epi_age_bracket[:] = rng.integers(0, 10, n_agents)
    st.soc_rate[:] = 0.8 + 0.4 * rng.random(n_agents); st.res_cellid[:] = 150
    st.seir[:] = np.where(rng.random(n_agents) < 0.2, 3, 1)
    cell_type = (np.arange(n_cells) % 11).astype(np.int8)
    h.set_cells(cell_type, np.full(n_cells, -1, np.int16))
    h.set_static(st.age_bracket, st.epi_age_bracket, st.soc_rate, st.res_cellid)
    h.set_state(seir=st.seir, inf_type=st.inf_type, inf_sev=st.inf_sev)
    d = h.make_day(capi.MODE_PHILOX, 5, 0.0, seed=2)
    h.simulate_contact_network(d, None, None, None, None, n=n, mem=capi.MEM_BUILT)
    rows_built = h.get_contacts()[0]
    h.simulate_contact_network(d, a, c, s, e)
    rows_host = h.get_contacts()[0]
    assert len(rows_built) > 100 and np.array_equal(rows_built, rows_host)
    ctx = mo.make_ctx(mo.MODE_PHILOX, 5, 0.0, False, False, seed=2)
    orow = mo.contact_network(P, ctx, oc, oa, os_, oe, cell_type, np.full(n_cells, -1, np.int16), st)[0]
    assert np.array_equal(rows_built, orow)
    h.close()
