"""Known answers of the reference's Epidemiology.simulate_seir_state_transition (epidemiology.py:233-500) called the way
the itinerary stage calls it for the agents that start the simulation Infectious (itinerary.py:373-395): an exposure day
BEFORE day 1, so that `reset_states` (:243) decides the agent's states.  Produced by importing the UNMODIFIED reference in
the build container; every random draw of every call is recorded in the slot layout of the replay tape's exposure records
(ref_harness.py EXPO_W) so that the oracle and the CUDA path can replay it:
    python tests/golden/make_seed_kat.py  -> tests/golden/seed_kat.npz (committed)"""
import json
import os
import random
import sys
from types import SimpleNamespace

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden import ref_harness  # noqa: E402

ref_harness._install_stubs()
sys.path.insert(0, "/root/reference/simulator")
import epidemiology as epi  # noqa: E402
import util as rutil  # noqa: E402

EXPO_W = ref_harness.EXPO_W
NONE = -32768            # the "None" sentinel of the dense records (mtd_b200.h MTD_NONE)
ep = json.load(open("/root/reference/data/epidemiology.json"))
N = 1200
dp = SimpleNamespace(quarantine_enabled=False, testing_enabled=False, contact_tracing_enabled=False,
                     masks_hygiene_distancing_multiplier=0.0)
vu = SimpleNamespace(agents_vaccination_doses={}, agents_seir_state={}, agents_infection_type={}, agents_infection_severity={})
agents_epi = {}
e = epi.Epidemiology(ep, N, 0, 1.0, None, agents_epi, vu, {}, {}, {}, dp)
probs = e.susceptibility_progression_mortality_probs_by_age

# ---- recording shims (draw order inside one call is fixed by the code: see the walk below)
seq = {"u": [], "ln": [], "ch": []}
orig_random, orig_choice, orig_sln = random.random, np.random.choice, rutil.sample_log_normal


def rec_random():
    u = orig_random(); seq["u"].append(u); return u


def rec_choice(a, size=None, replace=True, p=None):
    v = orig_choice(a, size=size, replace=replace, p=p); seq["ch"].append(int(np.asarray(v).reshape(-1)[0])); return v


def rec_sln(mean, std, size, isInt=False, recomputeparams=True):
    v = orig_sln(mean, std, size, isInt, recomputeparams); seq["ln"].append(int(v)); return v


epi.random.random = rec_random
np.random.choice = rec_choice
rutil.sample_log_normal = rec_sln

rs = np.random.RandomState(20240611)
random.seed(77); np.random.seed(78)
timesteps = np.arange(144)
T = dict(EXP_INF=0, INF_SYMP=1, SYMP_SEV=2, SEV_CRI=3, CRI_DEATH=4, ASYMP_REC=5, MILD_REC=6, SEV_REC=7, CRI_REC=8, REC_SUS=9)

cur_day = np.zeros(N, np.int32); expo_day = np.zeros(N, np.int32); bracket = np.zeros(N, np.int8)
asym = np.ones(N); testing = np.zeros(N, np.uint8); quar = np.zeros(N, np.uint8)
in_test = np.full((N, 2), NONE, np.int32); in_tres = np.full((N, 2), NONE, np.int32); in_hosp = np.full((N, 3), NONE, np.int32)
draws = np.full((N, EXPO_W), np.nan)
out_state = np.zeros((N, 4), np.int32)            # seir, inf type, severity, recovered
out_nsched = np.zeros(N, np.int32); out_sched = np.zeros((N, 8, 3), np.int32)
out_test = np.full((N, 2), NONE, np.int32); out_tres = np.full((N, 2), NONE, np.int32); out_hosp = np.full((N, 3), NONE, np.int32)
out_quar_unchanged = np.ones(N, np.uint8)

for i in range(N):
    cur_day[i] = 1 if i % 5 else 3
    # mostly before day 1 (reset_states), far enough back that whole trajectories can lie in the past; some on/after day 1
    expo_day[i] = cur_day[i] - int(rs.randint(1, 40)) if i % 7 else int(rs.randint(1, cur_day[i] + 1))
    bracket[i] = rs.randint(0, 10) if i % 3 else 9          # the oldest bracket reaches the severe / critical branches
    asym[i] = 1.0 if i % 11 else 0.1
    testing[i] = 1 if i % 2 else 0
    quar[i] = 1 if i % 4 < 2 else 0
    if i % 13 == 0:
        in_test[i] = (6, 70); in_tres[i] = (8, 70)
    if i % 17 == 0:
        in_hosp[i] = (2, 50, 9)
    dp.testing_enabled, dp.quarantine_enabled = bool(testing[i]), bool(quar[i])
    lst = lambda a: None if a[0] == NONE else [int(x) for x in a]
    q0 = [4, 55, 18] if i % 19 == 0 else None
    agent = {"state_transition_by_day": None, "test_day": lst(in_test[i]), "test_result_day": lst(in_tres[i]),
             "hospitalisation_days": lst(in_hosp[i]), "quarantine_days": q0}
    agents_epi[i] = agent
    for k in seq:
        seq[k].clear()
    # the caller's lines, itinerary.py:380-391
    trans = agent["state_transition_by_day"]
    agent_quarantine_days = agent["quarantine_days"]
    if trans is None:
        trans = []
    trans, seir, it, sv, recovered = e.simulate_seir_state_transition(int(cur_day[i]), agent, i, int(expo_day[i]), timesteps, trans,
                                                                      int(bracket[i]), agent_quarantine_days, float(asym[i]))
    agent["state_transition_by_day"] = trans
    agent["quarantine_days"] = agent_quarantine_days
    out_quar_unchanged[i] = agent["quarantine_days"] == q0
    # ---- draws -> slots, walking the branches the call took
    u, ln, ch = iter(seq["u"]), iter(seq["ln"]), iter(seq["ch"])
    row = draws[i]
    b = int(bracket[i])
    row[0] = next(ch)
    row[1 + T["EXP_INF"]] = next(ln)
    rec = False
    row[11] = next(u)
    symptomatic = row[11] < probs[1][b] * asym[i]
    if symptomatic:
        row[1 + T["INF_SYMP"]] = next(ln)
        row[12] = next(u)
        if row[12] < probs[2][b]:
            row[1 + T["SYMP_SEV"]] = next(ln)
            row[13] = next(u)
            if row[13] < probs[3][b]:
                row[1 + T["SEV_CRI"]] = next(ln)
                row[14] = next(u)
                if row[14] < probs[4][b]:
                    row[1 + T["CRI_DEATH"]] = next(ln)
                else:
                    row[1 + T["CRI_REC"]] = next(ln); rec = True
            else:
                row[1 + T["SEV_REC"]] = next(ln); rec = True
        else:
            row[1 + T["MILD_REC"]] = next(ln); rec = True
    else:
        row[1 + T["ASYMP_REC"]] = next(ln); rec = True
    if rec:
        row[1 + T["REC_SUS"]] = next(ln)
    assert rec == bool(recovered)
    # schedule_test / schedule_quarantine only when the symptomatic day is not before day 0 (epidemiology.py:479)
    if symptomatic and expo_day[i] + row[1 + T["EXP_INF"]] + row[1 + T["INF_SYMP"]] > -1:
        row[15] = next(ch)
        if testing[i] and in_test[i, 0] == NONE:
            row[16] = next(u)
            if row[16] < e.testing_after_symptoms_probability:
                row[17] = next(ln); row[18] = next(ln)
        row[19] = next(ch)
        if quar[i]:
            row[20] = next(u)
    for it_ in (u, ln, ch):
        assert next(it_, None) is None, "draw walk out of step with the reference at case %d" % i
    out_state[i] = (int(seir), int(it), int(sv), int(bool(recovered)))
    out_nsched[i] = len(trans)
    for k, (d, t, ts) in enumerate(trans):
        out_sched[i, k] = (int(d), int(t), int(ts))
    for name, arr in (("test_day", out_test), ("test_result_day", out_tres), ("hospitalisation_days", out_hosp)):
        v = agent[name]
        if v is not None and len(v) > 0:
            arr[i, :len(v)] = [int(x) for x in v]

assert out_quar_unchanged.all()
np.savez_compressed(os.path.join(HERE, "seed_kat.npz"), cur_day=cur_day, expo_day=expo_day, bracket=bracket, asym=asym,
                    testing=testing, quar=quar, in_test=in_test, in_tres=in_tres, in_hosp=in_hosp, draws=draws,
                    out_state=out_state, out_nsched=out_nsched, out_sched=out_sched, out_test=out_test, out_tres=out_tres,
                    out_hosp=out_hosp)
firsts = [int(out_sched[i, 0, 1]) if out_nsched[i] else -1 for i in range(N)]
print("first-kept transition histogram (reset cases):", {t: firsts.count(t) for t in sorted(set(firsts))})
print("states:", {tuple(s): int((out_state[:, :3] == s).all(1).sum()) for s in np.unique(out_state[:, :3], axis=0)})
print("tests scheduled:", int((out_test[:, 0] != NONE).sum() - (in_test[:, 0] != NONE).sum()), "hosp:", int((out_hosp[:, 0] != NONE).sum()))
