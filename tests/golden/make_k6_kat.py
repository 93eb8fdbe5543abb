"""Known answers of the intervention bookkeeping the reference's itinerary stage does once per agent and day
(Itinerary.sample_intervention_activities, itinerary.py:2283-2528): test results that are due today -> positive /
negative by the false-negative / false-positive rates -> contact_tracing_agent_ids + a Positive quarantine
(:2301-2332, schedule_quarantine epidemiology.py:568-625); expiry of test_result_day (:2331), hospitalisation_days on
their end day (:2396), quarantine_days on their end day (:2456), test_day six days after the test (:2486-2487).
Produced by running the UNMODIFIED reference (1k population, strict interventions) in the build container with the call
wrapped: records before / after, the SEIR state the call saw, the draws of random.random() it consumed, whether the
agent entered contact_tracing_agent_ids.

    python tests/golden/make_k6_kat.py [seeds...]   -> tests/golden/k6_kat.npz (committed)

Only calls of agents that are not travelling that day are kept (arrival / departure days re-time the records from the
itinerary, :2342-2383, which is outside this path); the number of dropped calls is printed.
"""
from __future__ import annotations

import os
import random
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
NONE = -32768
DAYS = 45


def flat(v, n):
    out = [NONE] * n
    if v is not None and len(v) > 0:
        for i, x in enumerate(v[:n]):
            out[i] = int(x)
    return out


def one(seed, out_path):
    import ref_harness as rh
    calls = []
    dropped = [0]

    def extra(rec):
        import itinerary as it
        orig = it.Itinerary.sample_intervention_activities
        state = {"draws": None}
        inner_random = random.random             # the harness's recording wrapper

        def rec_random():
            u = inner_random()
            if state["draws"] is not None:
                state["draws"].append(u)
            return u

        random.random = rec_random
        it.random.random = rec_random
        import epidemiology as epi
        epi.random.random = rec_random

        def wrapped(self, agentid, agent, agent_epi, agentindex, res_cellid, day, wakeup_ts, sleep_ts=None,
                    start_work_school_ts=None, end_work_school_ts=None, work_school_cellid=None, work_school_action=None,
                    is_departure_day_today=False, is_arrival_day_today=False, arr_dep_ts=None,
                    currently_on_travel_vacation=False, is_tourist=False, age_bracket_index=None, min_start_sleep_ts=None):
            keys = ("test_result_day", "test_day", "quarantine_days", "hospitalisation_days")
            busy = any(agent_epi.get(k) is not None and len(agent_epi[k]) > 0 for k in keys)
            if not busy:
                return orig(self, agentid, agent, agent_epi, agentindex, res_cellid, day, wakeup_ts, sleep_ts,
                            start_work_school_ts, end_work_school_ts, work_school_cellid, work_school_action,
                            is_departure_day_today, is_arrival_day_today, arr_dep_ts, currently_on_travel_vacation,
                            is_tourist, age_bracket_index, min_start_sleep_ts)
            pre = [flat(agent_epi["test_result_day"], 2), flat(agent_epi["test_day"], 2), flat(agent_epi["quarantine_days"], 3),
                   flat(agent_epi["hospitalisation_days"], 3)]
            seir = int(self.vars_util.agents_seir_state[agentid])
            was_traced = agentid in self.vars_util.contact_tracing_agent_ids
            state["draws"] = []
            try:
                out = orig(self, agentid, agent, agent_epi, agentindex, res_cellid, day, wakeup_ts, sleep_ts,
                           start_work_school_ts, end_work_school_ts, work_school_cellid, work_school_action,
                           is_departure_day_today, is_arrival_day_today, arr_dep_ts, currently_on_travel_vacation,
                           is_tourist, age_bracket_index, min_start_sleep_ts)
            finally:
                draws = state["draws"]; state["draws"] = None
            ae = out[1]
            post = [flat(ae["test_result_day"], 2), flat(ae["test_day"], 2), flat(ae["quarantine_days"], 3),
                    flat(ae["hospitalisation_days"], 3)]
            travelling = bool(is_departure_day_today or is_arrival_day_today or currently_on_travel_vacation)
            if travelling:
                dropped[0] += 1
                return out
            assert len(draws) <= 2
            calls.append(dict(day=int(day), agent=int(agentid), seir=seir, tourist=int(bool(is_tourist)),
                              qen=int(bool(self.epi_util.dyn_params.quarantine_enabled)), pre=pre, post=post,
                              draws=draws + [np.nan] * (2 - len(draws)),
                              traced=int((agentid in self.vars_util.contact_tracing_agent_ids) and not was_traced),
                              was_traced=int(was_traced)))
            return out

        it.Itinerary.sample_intervention_activities = wrapped

    wd = rh.prepare_workdir("/tmp/refwork_k6_%d" % seed, epidemiology_variant="epidemiology_strict_days.json")
    rh.run_reference(wd, seed=seed, days=DAYS, record_days=set(), extra_install=extra)
    n = len(calls)
    out = {"day_agent": np.array([[c["day"], c["agent"]] for c in calls], np.int32).reshape(-1, 2),
           "seir": np.array([c["seir"] for c in calls], np.int8),
           "flags": np.array([[c["tourist"], c["qen"], c["traced"], c["was_traced"]] for c in calls], np.int8).reshape(-1, 4),
           "draws": np.array([c["draws"] for c in calls], np.float64).reshape(-1, 2)}
    for k, (name, w) in enumerate((("tres", 2), ("test", 2), ("quar", 3), ("hosp", 3))):
        out["pre_" + name] = np.array([c["pre"][k] for c in calls], np.int32).reshape(-1, w)
        out["post_" + name] = np.array([c["post"][k] for c in calls], np.int32).reshape(-1, w)
    out["dropped"] = np.array([dropped[0]], np.int64)
    np.savez_compressed(out_path, **out)
    sys.stderr.write("seed %d: %d calls kept, %d dropped (travel days), %d results due, %d traced\n" % (
        seed, n, dropped[0], int(sum(c["pre"][0][0] == c["day"] for c in calls)), int(sum(c["traced"] for c in calls))))


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--one":
        one(int(sys.argv[2]), sys.argv[3])
        return
    seeds = [int(x) for x in sys.argv[1:]] or [21, 22, 23]
    procs = [(s, subprocess.Popen([sys.executable, os.path.abspath(__file__), "--one", str(s), "/tmp/k6_kat_%d.npz" % s],
                                  stdout=subprocess.DEVNULL)) for s in seeds]
    merged = {}
    for s, p in procs:
        if p.wait() != 0:
            raise SystemExit("seed %d failed" % s)
        z = np.load("/tmp/k6_kat_%d.npz" % s)
        for k in z.files:
            merged.setdefault(k, []).append(z[k])
    out = {k: np.concatenate(v) for k, v in merged.items()}
    np.savez_compressed(os.path.join(HERE, "k6_kat.npz"), **out)
    print("k6_kat.npz:", len(out["seir"]), "calls,", int(out["dropped"].sum()), "dropped")


if __name__ == "__main__":
    main()
