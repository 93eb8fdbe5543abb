"""Known answers of the reference's Itinerary.update_cell_agents_timesteps (itinerary.py:1403-1488) on 400 random
itineraries (runs of equal cells, -1 = residence, groups of agents, zero-length spans), produced by calling the
UNMODIFIED reference method in the build container:
    python tests/golden/make_intervals_kat.py  -> tests/golden/intervals_kat.npz (committed)"""
import contextlib
import io
import os
import sys
from types import SimpleNamespace

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden import ref_harness  # noqa: E402

ref_harness._install_stubs()
sys.path.insert(0, "/root/reference/simulator")
from itinerary import Itinerary  # noqa: E402

rng = np.random.default_rng(77)
it = Itinerary.__new__(Itinerary)
it.vars_util = SimpleNamespace(cells_agents_timesteps={})
ent_off, ent_ts, ent_cell, grp_off, grp_agent, grp_res = [0], [], [], [0], [], []
next_agent = 0
for t in range(400):
    n_ent = int(rng.integers(1, 9))
    ts = np.sort(rng.choice(np.arange(0, 144), size=n_ent, replace=False))
    if rng.random() < 0.3:
        ts[0] = 0
    cells = rng.choice(np.array([-1, -1, 3, 3, 7, 11, 12, 40]), size=n_ent)
    gsz = 1 if rng.random() < 0.7 else int(rng.integers(2, 5))
    agents = list(range(next_agent, next_agent + gsz)); next_agent += gsz
    res = [int(rng.integers(100, 120)) for _ in agents]
    itinerary = {int(a): (0, int(c)) for a, c in zip(ts, cells)}
    with contextlib.redirect_stdout(io.StringIO()):
        it.update_cell_agents_timesteps(5, itinerary, agents, res)
    ent_ts += [int(x) for x in ts]; ent_cell += [int(c) for c in cells]; ent_off.append(len(ent_ts))
    grp_agent += agents; grp_res += res; grp_off.append(len(grp_agent))
d = it.vars_util.cells_agents_timesteps
cell_order = np.array(list(d.keys()), np.int32)
off = np.cumsum([0] + [len(v) for v in d.values()])
flat = np.array([x for v in d.values() for x in v], np.int32).reshape(-1, 3)
np.savez_compressed(os.path.join(HERE, "intervals_kat.npz"), ent_off=np.array(ent_off, np.int64),
                    ent_ts=np.array(ent_ts, np.int16), ent_cell=np.array(ent_cell, np.int32),
                    grp_off=np.array(grp_off, np.int64), grp_agent=np.array(grp_agent, np.int32),
                    grp_rescell=np.array(grp_res, np.int32), cell_order=cell_order, cell_off=off.astype(np.int64), cell_rows=flat)
print("itineraries", 400, "intervals", len(flat), "cells", len(cell_order))
