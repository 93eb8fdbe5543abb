"""Derive the itinerary bank used for the tiled synthetic populations from the golden days.

    python tests/golden/make_bank.py      # writes mtdcovabm_b200/data/itinerary_bank_1k.npz

The bank holds, for every recorded day of the unmodified reference's config-1 run (tests/golden/ref_1k_seed6.npz), the
interval stream at the ContactNetwork.simulate_contact_network boundary (contactnetwork.py:65) plus the static agent
attributes and cell types of that population.  bench.py / mtdcovabm_b200.synth tile it to 10k / 520k / 5M agents.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from tests.golden import golden_io as gio  # noqa: E402


def main():
    z = gio.load_run("ref_1k_seed6")
    out = {"cell_type": z["cell_type"].astype(np.int8), "cell_indid": z["cell_indid"].astype(np.int16)}
    days = [int(d) for d in z["days"]]
    N = None
    for k, d in enumerate(days):
        g = gio.load_day(z, d)
        if N is None:
            N = g.N
            age = np.zeros(N, np.int8); epi = np.zeros(N, np.int8); soc = np.ones(N, np.float64)
            res = np.full(N, -1, np.int32); present = np.zeros(N, bool)
        ids = g.act_ids
        age[ids] = g.pre.age_bracket[ids]; epi[ids] = g.pre.epi_age_bracket[ids]
        soc[ids] = g.pre.soc_rate[ids]; res[ids] = g.pre.res_cellid[ids]
        present |= (g.pre.seir != 0)
        out["d%d_iv_agent" % k] = g.iv_agent.astype(np.int32); out["d%d_iv_cell" % k] = g.iv_cell.astype(np.int32)
        out["d%d_iv_start" % k] = g.iv_start.astype(np.uint8); out["d%d_iv_end" % k] = g.iv_end.astype(np.uint8)
    out.update(days=np.array(days, np.int32), n_agents=np.array([N], np.int64), age_bracket=age, epi_age_bracket=epi,
               soc_rate=soc, res_cellid=res, present=present)
    path = os.path.join(ROOT, "mtdcovabm_b200", "data", "itinerary_bank_1k.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()
