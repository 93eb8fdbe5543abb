"""Known answers of the reference's Epidemiology.schedule_test / schedule_quarantine / schedule_hospitalisation
(epidemiology.py:502-660) on a battery of calls with the legacy global RNG streams seeded (random.seed / np.random.seed),
produced by importing the UNMODIFIED reference in the build container:
    python tests/golden/make_schedule_kat.py   -> tests/golden/schedule_kat.json (committed)
The host mirrors in mtdcovabm_b200/epidemiology.py must return the same values, leave the same records and consume the
same number of draws (checked through the RNG state after the battery)."""
import json
import os
import random
import sys
from types import SimpleNamespace

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def battery(make, QT):
    """the call sequence both sides run: -> list of [kind, return values, record after]"""
    rs = np.random.RandomState(5)
    out = []
    for variant in ("epidemiology.json", "epidemiology_strict_days.json"):
        for testing in (False, True):
            for quar in (False, True):
                dp = SimpleNamespace(testing_enabled=testing, quarantine_enabled=quar)
                e, agents_epi = make(variant, dp)
                random.seed(31); np.random.seed(32)
                for i in range(60):
                    agent = {"test_day": None if i % 4 else [int(rs.randint(1, 30)), 50], "test_result_day": None,
                             "quarantine_days": None if i % 3 else [int(rs.randint(1, 30)), 60, 40],
                             "hospitalisation_days": None if i % 5 else [int(rs.randint(1, 20)), 70, 25]}
                    agents_epi[i] = agent
                    qt = QT(i % 4)
                    day = int(rs.randint(1, 30)); ts = int(rs.randint(42, 121))
                    r = e.schedule_test(agent, i, day, ts, qt)
                    out.append(["test", [bool(r[0]), r[1], r[2]], dict(agent)])
                    if i % 2:
                        r = e.schedule_quarantine(i, day, ts, qt, agent=agent)
                    else:
                        r = e.schedule_quarantine(i, day, ts, qt)
                    out.append(["quar", [bool(r[0]), r[1]], dict(agent)])
                    e.schedule_hospitalisation(agent, i, [day, ts, day + int(rs.randint(1, 20))])
                    out.append(["hosp", [], dict(agent)])
                out.append(["rng", [random.random(), float(np.random.random())], {}])
    return json.loads(json.dumps(out, default=lambda o: int(o)))


def make_ref(variant, dp):
    import epidemiology as epi      # the reference's (simulator/ is put on sys.path by __main__ below)
    ep = json.load(open("/root/reference/data/" + variant))
    agents_epi = {}
    vu = SimpleNamespace(agents_vaccination_doses={}, agents_seir_state={}, agents_infection_type={}, agents_infection_severity={})
    return epi.Epidemiology(ep, 100, 0, 1.0, None, agents_epi, vu, {}, {}, {}, dp), agents_epi


if __name__ == "__main__":
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from tests.golden import ref_harness
    ref_harness._install_stubs()
    sys.path.insert(0, "/root/reference/simulator")
    from epidemiologyclasses import QuarantineType
    kat = battery(make_ref, QuarantineType)
    json.dump(kat, open(os.path.join(HERE, "schedule_kat.json"), "w"))
    print(len(kat), "records;", sum(1 for k in kat if k[0] == "test" and k[1][0]), "tests and",
          sum(1 for k in kat if k[0] == "quar" and k[1][0]), "quarantines scheduled")
