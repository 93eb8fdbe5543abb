"""Known answers of the reference's Epidemiology.schedule_vaccinations (epidemiology.py:911-976) over 6 consecutive days
on a 600-local toy population, produced by importing the UNMODIFIED reference in the build container:
    python tests/golden/make_vacc_kat.py  -> tests/golden/vacc_kat.json (committed)"""
import json
import os
import sys
from types import SimpleNamespace

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden import ref_harness  # noqa: E402

ref_harness._install_stubs()
sys.path.insert(0, "/root/reference/simulator")
from epidemiology import Epidemiology  # noqa: E402

ep = json.load(open("/root/reference/data/epidemiology.json"))
n_locals, n_tourists, ratio = 600, 50, 0.02
agents_epi = {a: {"vaccination_days": None} for a in range(n_locals + n_tourists)}
vu = SimpleNamespace(agents_vaccination_doses={}, agents_seir_state={}, agents_infection_type={}, agents_infection_severity={})
dp = SimpleNamespace(vaccination_propensity=0.0, last_vaccination_propensity=0.0, num_agents_to_be_vaccinated=0,
                     num_vaccinations_today=0, quarantine_enabled=False, testing_enabled=False, contact_tracing_enabled=False,
                     masks_hygiene_distancing_multiplier=0.0)
e = Epidemiology(ep, n_locals, n_tourists, ratio, None, agents_epi, vu, {}, {}, {}, dp)
days = []
for day, prop in zip(range(10, 16), (0.0, 0.3, 0.3, 0.5, 0.5, 0.5)):
    dp.vaccination_propensity = prop
    np.random.seed(1000 + day)          # the mirror draws its agents on the device: re-seed per call
    upd = e.schedule_vaccinations(day)
    ts = [int(agents_epi[a]["vaccination_days"][-1][1]) for a in upd]
    days.append({"day": day, "propensity": prop, "updated": [int(a) for a in upd], "ts": ts,
                 "num_agents_to_be_vaccinated": int(dp.num_agents_to_be_vaccinated),
                 "num_vaccinations_today": int(dp.num_vaccinations_today),
                 "last_vaccination_propensity": float(dp.last_vaccination_propensity)})
    # the itinerary stage of the next day administers the doses (itinerary.py:2517-2522)
    for a in upd:
        vu.agents_vaccination_doses.setdefault(a, []).append(day + 1)
json.dump({"n_locals": n_locals, "n_tourists": n_tourists, "ratio": ratio, "seed": 42, "days": days},
          open(os.path.join(HERE, "vacc_kat.json"), "w"))
print([(d["day"], d["num_vaccinations_today"], len(d["updated"]), d["num_agents_to_be_vaccinated"]) for d in days])
