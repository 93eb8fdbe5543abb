#!/usr/bin/env python
"""Cost of the reference-shaped boundary (SURVEY.md section 7 "report both kernel-only and through the Python boundary"):
times ContactNetwork.simulate_contact_network(day, weekday) of the drop-in class -- dicts of lists in, list of 5-int rows
and per-agent dicts out, exactly the reference's containers (contactnetwork.py:65-124) -- on the recorded golden days of
config 1 (1k agents) and of the tiled 10k population (configs 2 / 5), next to the device time of the same day through the
array API (capi.Handle), in Philox mode.  GPU box:   python tools/time_dict_api.py > profiles/r02_dict_api_cost.json"""
import json
import os
import sys
import time
from types import SimpleNamespace

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mtdcovabm_b200 import capi, params as prm  # noqa: E402
from mtdcovabm_b200.contactnetwork import ContactNetwork  # noqa: E402
from tests.golden import golden_io as gio  # noqa: E402
from tests.test_dropin_api import FakeStatic, reference_containers  # noqa: E402
from tests import helpers as H  # noqa: E402


def main():
    out = []
    for name in ("ref_1k_seed6", "ref_10k_normal_seed11", "ref_10k_strict_seed12"):
        z = gio.load_run(name)
        ep = prm.load_json(gio.RUNS[name])
        P = prm.build_params(None, ep)
        static = None                      # ONE Static object per run, as in the reference (static.py: built once by sim.py)
        for day in [int(d) for d in z["days"]]:
            g = gio.load_day(z, day)
            static = static or FakeStatic(g.pre)
            pairs = None
            t_dict = []
            for rep in range(3):
                epi, vu = reference_containers(g, g.pre)
                dyn = SimpleNamespace(masks_hygiene_distancing_multiplier=g.mhd, testing_enabled=g.testing_enabled,
                                      quarantine_enabled=g.quarantine_enabled, contact_tracing_enabled=g.contact_tracing_enabled)
                cells_type = {c: int(t) for c, t in enumerate(g.cell_type) if t >= 0}
                indids = {c: int(v) for c, v in enumerate(g.cell_indid) if v >= 0}
                cn = ContactNetwork(g.n_locals, g.N - g.n_locals, 1.0, static, epi, vu, cells_type, indids, {}, {}, {},
                                    prm.load_json("contactnetwork.json"), ep, dyn, mode="philox", seed=1)
                t0 = time.perf_counter()
                cn.simulate_contact_network(g.day, g.weekday)
                t_dict.append(time.perf_counter() - t0)
                n_rows = len(vu.directcontacts_by_simcelltype_by_day)
            # the same day through the array API: device time of K0..K3 with the inputs already uploaded
            h = H.make_handle(P, g.pre, g.cell_type, g.cell_indid, len(g.iv_cell) + 16)
            d = h.make_day(capi.MODE_PHILOX, g.day, g.mhd, g.testing_enabled, g.quarantine_enabled, seed=1)
            t_arr = []
            for rep in range(3):
                H.upload_state(h, g.pre)
                h.sync()
                t0 = time.perf_counter()
                h.simulate_contact_network(d, g.iv_agent, g.iv_cell, g.iv_start, g.iv_end)
                h.simulate_direct_contacts(d)
                c = h.counters()
                t_arr.append(time.perf_counter() - t0)
            pairs = int(c.pairs_evaluated)
            h.close()
            out.append({"run": name, "day": day, "intervals": int(len(g.iv_cell)), "pairs": pairs, "rows": n_rows,
                        "dict_api_ms": 1e3 * min(t_dict), "array_api_ms": 1e3 * min(t_arr),
                        "dict_api_pairs_per_s": pairs / min(t_dict), "array_api_pairs_per_s": pairs / min(t_arr)})
            sys.stderr.write("%s day %d: dict %.1f ms, arrays %.2f ms, %d pairs\n" % (name, day, 1e3 * min(t_dict), 1e3 * min(t_arr), pairs))
    summ = {}
    for name in ("ref_1k_seed6", "ref_10k_normal_seed11", "ref_10k_strict_seed12"):
        r = [o for o in out if o["run"] == name]
        summ[name] = {"days": len(r), "dict_api_pairs_per_s": sum(o["pairs"] for o in r) / (1e-3 * sum(o["dict_api_ms"] for o in r)),
                      "array_api_pairs_per_s": sum(o["pairs"] for o in r) / (1e-3 * sum(o["array_api_ms"] for o in r)),
                      "dict_api_ms_per_day": sum(o["dict_api_ms"] for o in r) / len(r),
                      "array_api_ms_per_day": sum(o["array_api_ms"] for o in r) / len(r)}
    print(json.dumps({"what": "ContactNetwork.simulate_contact_network (reference dict containers) vs the array API on the same recorded days, "
                              "host wall clock incl. uploads and the row download; the reference's own Python path runs these days at "
                              "1.5e5 (1k) / 1.3e5 (10k) pairs/s on one core (profiles/r02_reference_python_timing.json)",
                      "summary": summ, "days": out}, indent=1))


if __name__ == "__main__":
    main()
