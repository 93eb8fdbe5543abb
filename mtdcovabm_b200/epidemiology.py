"""Drop-in for the hot-path methods of the reference's `epidemiology.Epidemiology` (epidemiology.py:14-906): same
constructor; `contact_tracing(day, distributed=False, f=None)` (epidemiology.py:667-834) and
`contact_tracing_clean_up(day, f=None)` (:838-906) computed through libmtd_b200.so (K5).  The transmission step
(`simulate_direct_contacts`, :88-231) runs fused behind `ContactNetwork.simulate_contact_network` (K3); there is no
CPU fallback for any of it.

Extra keyword arguments (not in the reference): `mode` ("philox" | "replay"), `seed`, `tape`, `device`.
"""
from __future__ import annotations

import numpy as np

from . import params as prm


class Epidemiology:
    def __init__(self, epidemiologyparams, n_locals, n_tourists, locals_ratio_to_full_pop, agents_static, agents_epi,
                 vars_util, cells_households, cells_institutions, cells_accommodation, dyn_params, process_index=-1,
                 agents_seir_indices=None, *, mode="philox", seed=0, tape=None, device=0):
        self.n_locals, self.n_tourists = n_locals, n_tourists
        self.locals_ratio_to_full_pop = locals_ratio_to_full_pop
        self.epidemiologyparams = epidemiologyparams
        ep = epidemiologyparams
        self.contact_tracing_days_back, self.contact_tracing_positive_delay_days, \
            self.contact_tracing_secondary_delay_days = ep["contact_tracing_parameters"]
        self.dyn_params = dyn_params
        self.agents_static, self.agents_epi, self.vars_util = agents_static, agents_epi, vars_util
        self.agents_seir_indices = agents_seir_indices
        self.cells_households, self.cells_institutions = cells_households, cells_institutions
        self.cells_accommodation = cells_accommodation
        self.timestep_options = np.arange(42, 121)          # 7.00 - 20.00 (epidemiology.py:84)
        self.process_index = process_index
        self.mode, self.seed, self.tape, self.device = mode, seed, tape, device
        self._params = None

    @property
    def params(self):
        if self._params is None:
            self._params = prm.build_params(None, self.epidemiologyparams)
        return self._params

    def simulate_direct_contacts(self, agents_directcontacts, cellid, cell_type, day):
        """epidemiology.py:88-231.  On the B200 path this step is fused behind ContactNetwork.simulate_contact_network
        (K3 consumes the contact rows K1/K2 left in HBM); a per-cell Python dict of contacts never exists."""
        raise RuntimeError("simulate_direct_contacts runs fused inside ContactNetwork.simulate_contact_network on the "
                           "B200 path (mtd_simulate_direct_contacts); call that instead")
