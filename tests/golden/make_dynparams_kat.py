"""Known answers of the reference's DynamicParams.refresh_dynamic_parameters (dynamicparams.py:59-134) over the three
epidemiology parameter files the configs use, produced by importing the UNMODIFIED reference in the build container:
    python tests/golden/make_dynparams_kat.py   -> tests/golden/dynparams_kat.json (committed)"""
import io
import contextlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden import ref_harness  # noqa: E402

ref_harness._install_stubs()
sys.path.insert(0, "/root/reference/simulator")
from dynamicparams import DynamicParams  # noqa: E402

FIELDS = ["masks_hygiene_distancing_multiplier", "vaccination_propensity", "quarantine_enabled", "testing_enabled",
          "contact_tracing_enabled", "travel_restrictions_multiplier", "workplaces_lockdown", "schools_lockdown",
          "entertainment_lockdown", "activities_lockdown", "airport_lockdown"]
out = {}
for name in ("epidemiology.json", "epidemiology_normal_days.json", "epidemiology_strict_days.json"):
    ep = json.load(open(os.path.join("/root/reference/data", name)))
    variants = {"day": ep}
    # a rate-threshold variant of the same file (the reference prefers rate thresholds when they are given)
    er = dict(ep)
    er["masks_hygiene_distancing_infectiousrate_thresholds"] = [[0.0, 0.0], [0.01, 0.2], [0.05, 0.5]]
    er["intervention_infectiousrate_thresholds"] = [0.005, 0.01, 0.02]
    er["lockdown_infectiousrate_thresholds"] = [0.02, 0.03, 0.04, 0.05, 0.06]
    er["vaccination_infectiousrate_thresholds"] = [[0.0, 0.0], [0.02, 0.1]]
    er["travel_restrictions_infectiousrate_thresholds"] = [[0.0, 0.0], [0.03, 0.6]]
    variants["rate"] = er
    for vn, e in variants.items():
        dp = DynamicParams(1000, 2000, 40, e)
        rows = []
        for day in range(1, 61):
            inf = int(3 + 2.5 * day if day < 30 else max(0, 78 - 3 * (day - 30)))
            dec = day // 10
            seir = (dec, 5 + day, 1000 - inf - dec - 5 - day, inf, 0)      # deceased, exposed, susceptible, infectious, recovered
            with contextlib.redirect_stdout(io.StringIO()):
                dp.statistics.calculate_seir_states_counts = lambda vu, a=False, b=False, s=seir: s
                dp.refresh_dynamic_parameters(day, 40, 3, 2, 1, None, True)
            rows.append({"day": day, "seir": list(seir), "infectious_rate": dp.statistics.infectious_rate,
                         "new_infectious": dp.statistics.new_infectious,
                         **{f: (float(getattr(dp, f)) if not isinstance(getattr(dp, f), bool) else bool(getattr(dp, f)))
                            for f in FIELDS}})
        out[name + ":" + vn] = {"params": {k: e[k] for k in e if k.endswith("_thresholds")}, "rows": rows}
json.dump(out, open(os.path.join(HERE, "dynparams_kat.json"), "w"))
print("wrote dynparams_kat.json:", {k: len(v["rows"]) for k, v in out.items()})
