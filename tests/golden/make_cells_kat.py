"""Known answers of the reference's ContactNetwork on HAND-MADE cells (SURVEY.md section 8c pin (2)) and of
seirstateutil.initialize_agent_states, produced by importing the UNMODIFIED reference in the build container:

    python tests/golden/make_cells_kat.py      -> tests/golden/ref_cells_kat.npz, tests/golden/init_states_kat.npz

One synthetic "day" whose cells are the edge cases of generate_contact_network (contactnetwork.py:134-315): a lone
interval, one agent twice, touching intervals (end == start), disjoint intervals, a repeated agent with a partner
(first start / last end of the overlap list, :101), a Workplace cell with indid == 9 (-> activity column 5, :317-347), a
Classroom cell (-> column 0), a Workplace cell where an agent lives (base probability quirk, epidemiology.py:998-1014),
plus random cells of 40 .. 1100 intervals (one for every size class of the block-per-cell K1 kernel, agents that come
back several times included).  It goes through ContactNetwork.simulate_contact_network under the same recorder as the
whole-run fixtures (ref_harness.install_recorder), so the file has the format of the other golden days: inputs, contact
rows, exposed ids, post state and the keyed draw tape.  A third of the agents is Infectious, so the transmission stage
(epidemiology.py:88-231) runs on the hand-made rows as well.
"""
from __future__ import annotations

import json
import os
import random
import sys
from types import SimpleNamespace

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness as rh  # noqa: E402
import make_golden  # noqa: E402

rh._install_stubs()
sys.path.insert(0, "/root/reference/simulator")
import contactnetwork as cn  # noqa: E402
import seirstateutil as ssu  # noqa: E402
from epidemiologyclasses import SEIRState, InfectionType, Severity  # noqa: E402

REF = "/root/reference"


class StaticStub:
    """Static.get(id, name) (static.py:351-363) over plain arrays"""

    def __init__(self, **arrs):
        self.a = arrs

    def get(self, agentid, name):
        return self.a[name][agentid]


def build_cells(rng):
    """-> {cellid: [(agent, start, end), ...]} (dict order = cell order), cell types, indids, number of agents"""
    cells, ctype, indid = {}, {}, {}
    nxt = [0]

    def fresh(k):
        a = list(range(nxt[0], nxt[0] + k)); nxt[0] += k
        return a

    def add(ct, ivs, ind=None):
        cid = len(cells)
        cells[cid] = [(int(a), int(s), int(e)) for a, s, e in ivs]
        ctype[cid] = ct
        if ind is not None:
            indid[cid] = ind
        return cid

    a = fresh(1); add(0, [(a[0], 0, 143)])                                           # n = 1
    a = fresh(1); add(0, [(a[0], 0, 20), (a[0], 50, 70)])                            # one agent twice
    a = fresh(2); add(0, [(a[0], 10, 40), (a[1], 40, 80)])                           # touching: one shared timestep
    a = fresh(2); add(0, [(a[0], 10, 40), (a[1], 41, 80)])                           # disjoint
    a = fresh(2); add(0, [(a[0], 0, 10), (a[1], 20, 30), (a[0], 25, 40), (a[1], 5, 8)])   # SURVEY Appendix A probe
    a = fresh(3); add(1, [(a[0], 0, 30), (a[1], 20, 70), (a[0], 60, 100), (a[2], 0, 143)], ind=9)   # indid 9 -> column 5
    a = fresh(4); add(1, [(a[0], 0, 60), (a[1], 30, 90), (a[2], 10, 50), (a[3], 45, 46)], ind=3)    # plain workplace
    a = fresh(6); add(6, [(x, 5 * i, 60 + 5 * i) for i, x in enumerate(a)])          # Classroom -> "total" column
    a = fresh(5); add(5, [(x, 48, 96) for x in a])                                   # School
    a = fresh(5); add(8, [(x, 40 + i, 44 + i) for i, x in enumerate(a)])             # Transport
    a = fresh(7); add(10, [(x, 30, 70 + i) for i, x in enumerate(a)] + [(a[0], 100, 120), (a[3], 90, 105)])   # Airport
    for m in (12, 40, 100, 170, 260, 420, 1100):                                     # random cells, every K1 size class
        k = max(2, int(m * 0.8))
        a = np.array(fresh(k))
        ag = a[rng.integers(0, k, m)]
        ag[:k] = rng.permutation(a)                                                  # everybody at least once
        s = rng.integers(0, 120, m); e = np.minimum(143, s + rng.integers(0, 50, m))
        add(int(rng.choice([0, 1, 4, 10, 8, 3])), zip(ag, s, e), ind=None)
    return cells, ctype, indid, nxt[0]


def main():
    random.seed(12); np.random.seed(13)
    rng = np.random.default_rng(14)
    cells, ctype, indid, N = build_cells(rng)
    ep = json.load(open(os.path.join(REF, "data", "epidemiology.json")))
    cnp = json.load(open(os.path.join(REF, "data", "contactnetwork.json")))
    static = StaticStub(age_bracket_index=rng.integers(0, 16, N), epi_age_bracket_index=rng.integers(0, 10, N),
                        soc_rate=0.8 + 0.4 * rng.random(N), res_cellid=rng.integers(0, len(cells), N))
    static.a["res_cellid"][13] = 6                                                   # lives in the plain workplace cell
    seir = {i: (SEIRState.Infectious if rng.random() < 0.33 else SEIRState.Susceptible) for i in range(N)}
    vu = SimpleNamespace(cells_agents_timesteps=cells, directcontacts_by_simcelltype_by_day=[],
                         directcontacts_by_simcelltype_by_day_start_marker={}, agents_seir_state=seir,
                         agents_infection_type={i: InfectionType.Asymptomatic for i in range(N) if seir[i] == SEIRState.Infectious},
                         agents_infection_severity={i: Severity.Mild for i in range(N) if seir[i] == SEIRState.Infectious},
                         agents_vaccination_doses={}, contact_tracing_agent_ids=set())
    agents_epi = {i: {"state_transition_by_day": None, "test_day": None, "test_result_day": None,
                      "hospitalisation_days": None, "quarantine_days": None, "vaccination_days": None} for i in range(N)}
    dyn = SimpleNamespace(masks_hygiene_distancing_multiplier=0.1, testing_enabled=True, quarantine_enabled=True,
                          contact_tracing_enabled=False)
    rec = rh.Recorder()
    rh.install_recorder(rec)
    net = cn.ContactNetwork(N, 0, 1.0, static, agents_epi, vu, ctype, indid, {}, {}, {}, cnp, ep, dyn)
    net.simulate_contact_network(1, 1)
    assert len(rec.days) == 1
    d = rec.days[0]
    out = make_golden.encode_run(rec, [1])
    np.savez_compressed(os.path.join(HERE, "ref_cells_kat.npz"), **out)
    rows = d["contacts"]
    print("cells:", len(cells), "intervals:", len(d["iv_cell"]), "rows:", len(rows), "exposed:", len(d["exposed_ids"]),
          "selections:", len(d["tape_sel"]))
    # the documented known answers
    by_pair = {(int(r[1]), int(r[2])): (int(r[3]), int(r[4])) for r in rows}
    print("touching:", by_pair.get((2, 3)), " appendix-A probe:", by_pair.get((6, 7)))

    # ---- seirstateutil.initialize_agent_states (seirstateutil.py:5-32), legacy global RNG streams seeded
    kat = {}
    for k, (n, dist, s1, s2) in enumerate([(1000, [0.95, 0.0, 0.05, 0.0, 0.0], 3, 4), (1015, [0.9, 0.02, 0.05, 0.03, 0.0], 5, 6),
                                           (37, [0.5, 0.1, 0.1, 0.1, 0.0], 7, 8), (10, [0.33, 0.33, 0.0, 0.0, 0.0], 9, 10)]):
        np.random.seed(s1); random.seed(s2)
        st = {i: SEIRState.Undefined for i in range(n)}
        st = ssu.initialize_agent_states(n, dist, st)
        kat["c%d_n" % k] = np.array([n, s1, s2]); kat["c%d_dist" % k] = np.array(dist)
        kat["c%d_out" % k] = np.array([int(st[i]) for i in range(n)], np.int8)
    np.savez_compressed(os.path.join(HERE, "init_states_kat.npz"), **kat)
    print("init states:", {k: np.bincount(v, minlength=6).tolist() for k, v in kat.items() if k.endswith("_out")})


if __name__ == "__main__":
    main()
