"""Known-answer vectors of the reference's hot-path helpers (simulator/util.py:88-150, 239-297;
seirstateutil.py:105-136; contactnetwork.py:317-347; epidemiology.py:998-1014), produced by importing the UNMODIFIED
reference from /root/reference in the build container:   python tests/golden/make_util_kat.py
Writes tests/golden/util_kat.npz (committed; /root/reference does not exist on the GPU box)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden import ref_harness  # noqa: E402

ref_harness._install_stubs()
sys.path.insert(0, "/root/reference/simulator")
import util as rutil  # noqa: E402
import seirstateutil as rseir  # noqa: E402
from cellsclasses import CellType  # noqa: E402
from epidemiologyclasses import SEIRState, InfectionType, Severity, SEIRStateTransition  # noqa: E402

out = {}
# sample_log_normal: parameter pairs of data/epidemiology.json durations, fixed legacy seed
pars = [(4.5, 1.5), (1.1, 0.9), (6.6, 4.9), (1.5, 2.0), (10.7, 4.8), (8.0, 2.0), (18.1, 6.3), (90.0, 15.0), (1.0, 0.5)]
draws = []
for k, (m, s) in enumerate(pars):
    np.random.seed(100 + k)
    draws.append([rutil.sample_log_normal(m, s, 1, True) for _ in range(8)])
np.random.seed(7)
out["ln_pars"] = np.array(pars)
out["ln_int_draws"] = np.array(draws, np.int64)
np.random.seed(7)
out["ln_float_draws"] = np.array(rutil.sample_log_normal(4.5, 1.5, 16, False))
# vaccination multipliers over days since the dose
vm = []
for d in range(0, 401):
    i, a, _ = rutil.calculate_vaccination_multipliers({5: [10]}, 5, 10 + d, 0.9, 0.8, 30)
    vm.append((i, a))
out["vacc_mult"] = np.array(vm)
out["vacc_none"] = np.array(rutil.calculate_vaccination_multipliers({5: [10]}, 6, 50, 0.9, 0.8, 30)[:2])
# cell type tables
out["simcelltype"] = np.array([int(rutil.convert_celltype_to_simcelltype(0, celltype=CellType(t))) for t in range(11)])
# partner lists
keys = [(1, 2), (3, 1), (2, 3), (4, 1), (2, 4), (5, 6)]
ids, deg = [1, 2, 3, 4, 5, 6], [2, 0, 1, 3, 1, 0]
out["partners_plain"] = np.array([list(rutil.get_all_contacts_ids_by_id(1, keys)) + [-1] * 0])
out["partners_alive"] = np.array(list(rutil.get_all_contacts_ids_by_id(1, keys, ids, deg)))
np.random.seed(11)
out["partners_shuffled"] = np.array(list(rutil.get_all_contacts_ids_by_id(1, keys, ids, [1] * 6, shuffle=True)))
out["partners_none"] = np.array([rutil.get_all_contacts_ids_by_id(5, keys, ids, deg) is None])
# transition table over every input (seirstateutil.py:105-136)
tab = []
for s in range(6):
    for it in range(5):
        for sv in range(4):
            for tr in range(10):
                r = rseir.convert_state_transition_to_new_state(SEIRState(s), InfectionType(it), Severity(sv), SEIRStateTransition(tr))
                tab.append([s, it, sv, tr, int(r[0]), int(r[1]), int(r[2])])
out["transition_table"] = np.array(tab, np.int32)
np.savez_compressed(os.path.join(HERE, "util_kat.npz"), **out)
print("wrote util_kat.npz:", {k: v.shape for k, v in out.items()})
