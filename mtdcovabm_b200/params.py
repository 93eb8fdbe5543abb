"""Model parameters for the hot path, read from the reference's JSON parameter files.

Mirrors what `ContactNetwork.__init__` (contactnetwork.py:41-42) and `Epidemiology.__init__`
(epidemiology.py:29-70) pull out of `contactnetworkparams` / `epidemiologyparams`, flattened into the
`mtd_params` C struct declared in include/mtd_b200.h.  All derived constants (k per activity, log-normal
underlying-normal parameters) are computed here with the *same numpy expressions the reference uses*, so that the
device never has to reproduce numpy's rounding of those.
"""
from __future__ import annotations

import ctypes as C
import json
import os

import numpy as np

DATA_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")

MAXSCHED = 12
NONE = -32768
EXPO_W = 21


class ParamsStruct(C.Structure):
    """Binary layout of `mtd_params` (include/mtd_b200.h)."""
    _fields_ = [
        ("cm", C.c_double * 6 * 16),
        ("k_by_act", C.c_int32 * 6),
        ("p_household", C.c_double), ("p_workplace", C.c_double), ("p_school", C.c_double), ("p_community", C.c_double),
        ("probs_by_age", C.c_double * 11 * 5),
        ("ln_mu", C.c_double * 10), ("ln_sigma", C.c_double * 10),
        ("test_mu", C.c_double), ("test_sigma", C.c_double), ("result_mu", C.c_double), ("result_sigma", C.c_double),
        ("testing_after_symptoms_probability", C.c_double),
        ("quarantine_after_symptoms_probability", C.c_double),
        ("ignore_quarantine_rules_probability", C.c_double),
        ("quarantine_positive_duration", C.c_int32), ("quarantine_positive_contact_duration", C.c_int32),
        ("quarantine_secondary_contact_duration", C.c_int32),
        ("contact_tracing_days_back", C.c_int32), ("contact_tracing_positive_delay_days", C.c_int32),
        ("contact_tracing_secondary_delay_days", C.c_int32),
        ("tracing_success_prob", C.c_double * 4),
        ("vacc_immunity", C.c_double), ("vacc_asymptomatic", C.c_double), ("vacc_decay_interval", C.c_double),
        ("testing_false_positive_rate", C.c_double), ("testing_false_negative_rate", C.c_double),
    ]


def lognormal_params(mean, std):
    """util.sample_log_normal's parameter conversion (util.py:89-91), same numpy calls."""
    mu = np.log(mean ** 2 / np.sqrt(std ** 2 + mean ** 2))
    sigma = np.sqrt(np.log(std ** 2 / mean ** 2 + 1))
    return float(mu), float(sigma)


def load_json(name_or_path):
    path = name_or_path if os.path.isabs(name_or_path) or os.path.exists(name_or_path) \
        else os.path.join(DATA_DIR, name_or_path)
    with open(path) as f:
        return json.load(f)


def build_params(contactnetworkparams=None, epidemiologyparams=None) -> ParamsStruct:
    """dicts (as loaded by sim.py:410-459) -> mtd_params."""
    cnp = contactnetworkparams if contactnetworkparams is not None else load_json("contactnetwork.json")
    ep = epidemiologyparams if epidemiologyparams is not None else load_json("epidemiology.json")
    P = ParamsStruct()
    mat = np.array(cnp["ageactivitycontactmatrix"])                      # contactnetwork.py:42
    for r in range(16):
        for a in range(6):
            P.cm[r][a] = float(mat[r, 2 + a])
    for a in range(6):
        P.k_by_act[a] = int(round(np.mean(np.array(mat[:, 2 + a]))))     # contactnetwork.py:212-214
    P.p_household = ep["household_infection_probability"]
    P.p_workplace = ep["workplace_infection_probability"]
    P.p_school = ep["school_infection_probability"]
    P.p_community = ep["community_infection_probability"]
    tab = ep["susceptibility_progression_mortality_probs_by_age"]
    for r in range(5):
        for c in range(11):
            P.probs_by_age[r][c] = float(tab[r][c])
    for t, (mean, std) in enumerate(ep["durations_days_distribution_parameters"]):   # epidemiology.py:38-51
        P.ln_mu[t], P.ln_sigma[t] = lognormal_params(mean, std)
    P.test_mu, P.test_sigma = lognormal_params(*ep["testing_days_distribution_parameters"])
    P.result_mu, P.result_sigma = lognormal_params(*ep["testing_results_days_distribution_parameters"])
    P.testing_after_symptoms_probability = ep["testing_after_symptoms_probability"]
    P.quarantine_after_symptoms_probability = ep["quarantine_after_symptoms_probability"]
    P.ignore_quarantine_rules_probability = ep["ignore_quarantine_rules_probability"]
    q = ep["quarantine_duration_parameters"]
    P.quarantine_positive_duration, P.quarantine_positive_contact_duration, P.quarantine_secondary_contact_duration = q
    ct = ep["contact_tracing_parameters"]
    P.contact_tracing_days_back, P.contact_tracing_positive_delay_days, P.contact_tracing_secondary_delay_days = ct
    for i, v in enumerate(ep["contact_tracing_success_probability"]):
        P.tracing_success_prob[i] = v
    v = ep["vaccination_effectiveness_parameters"]
    P.vacc_immunity, P.vacc_asymptomatic, P.vacc_decay_interval = float(v[0]), float(v[1]), float(v[2])
    fp, fn = ep["testing_false_positive_negative_rates"]                    # epidemiology.py:59-60
    P.testing_false_positive_rate, P.testing_false_negative_rate = float(fp), float(fn)
    return P
