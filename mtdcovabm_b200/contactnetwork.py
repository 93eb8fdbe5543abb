"""Drop-in for the reference's `contactnetwork.ContactNetwork` (contactnetwork.py:11-124): same constructor, same
`simulate_contact_network(day, weekday)` contract, computed by libmtd_b200.so (K0 CSR build, K1 per-cell contact
network, K2 rows, K3 transmission) on a B200.  There is no CPU fallback.

Extra keyword arguments (not in the reference): `mode` ("philox" | "replay"), `seed` (Philox key), `tape` (replay
tape: the reference's own draws, see tests/golden/ref_harness.py) and `device`.
"""
from __future__ import annotations

import time

import numpy as np

from . import bridge, capi, params as prm
from .epidemiology import Epidemiology


class ContactNetwork:
    def __init__(self, n_locals, n_tourists, locals_ratio_to_full_pop, agents_static, agents_epi, vars_util, cells_type,
                 indids_by_cellid, cells_households, cells_institutions, cells_accommodation, contactnetworkparams,
                 epidemiologyparams, dynparams, contact_network_sum_time_taken=0, visualise=False,
                 maintain_directcontacts_count=False, process_index=-1, *, mode="philox", seed=0, tape=None, device=0):
        self.n_locals, self.n_tourists = n_locals, n_tourists
        self.locals_ratio_to_full_pop = locals_ratio_to_full_pop
        self.agents_static, self.agents_epi, self.vars_util = agents_static, agents_epi, vars_util
        self.cells_type, self.indids_by_cellid = cells_type, indids_by_cellid
        self.cells_households, self.cells_institutions = cells_households, cells_institutions
        self.cells_accommodation = cells_accommodation
        self.contactnetworkparams, self.epidemiologyparams = contactnetworkparams, epidemiologyparams
        self.dyn_params = dynparams
        self.contactnetwork_sum_time_taken = contact_network_sum_time_taken
        self.visualise, self.maintain_directcontacts_count = visualise, maintain_directcontacts_count
        self.process_index = process_index
        self.mode, self.seed, self.tape, self.device = mode, seed, tape, device
        self.params = prm.build_params(contactnetworkparams, epidemiologyparams)
        # the reference builds its Epidemiology helper here too (contactnetwork.py:62)
        self.epi_util = Epidemiology(epidemiologyparams, n_locals, n_tourists, locals_ratio_to_full_pop, agents_static,
                                     agents_epi, vars_util, cells_households, cells_institutions, cells_accommodation,
                                     dynparams, process_index, mode=mode, seed=seed, device=device)
        self.epi_util.cn_tape = tape
        self.counters = None

    # ---- the reference's cell tables -> dense arrays indexed by cell id
    def _cell_tables(self, max_cell):
        ct = np.full(max_cell + 1, -1, np.int8)
        ci = np.full(max_cell + 1, -1, np.int16)
        for c, t in self.cells_type.items():
            if 0 <= int(c) <= max_cell:
                ct[int(c)] = int(t)
        for c, v in self.indids_by_cellid.items():
            if v is not None and 0 <= int(c) <= max_cell:
                ci[int(c)] = int(v)
        return ct, ci

    # ---- single-cell entry points of the reference (contactnetwork.py:127-315), same device path on one cell
    def _run_cell(self, cellid, day, transmit):
        """K0-K2 (and K3 when `transmit`) over the interval list of one cell -> (contacts dict, exposed ids)"""
        vu = self.vars_util
        lst = vu.cells_agents_timesteps[cellid]
        M = len(lst)
        contacts, exposed = {}, []
        if M == 0:
            return contacts, exposed
        iv_cell = np.full(M, int(cellid), np.int32)
        iv_agent = np.array([int(a) for a, _, _ in lst], np.int32)
        iv_start = np.array([int(s) for _, s, _ in lst], np.uint8)
        iv_end = np.array([int(e) for _, _, e in lst], np.uint8)
        n_agents = int(self.n_locals + self.n_tourists)
        max_cell = max(int(cellid), max((int(c) for c in self.cells_type.keys()), default=0))
        cell_type, cell_indid = self._cell_tables(max_cell)
        dense = bridge.flatten(n_agents, self.agents_static, self.agents_epi, vu, np.unique(iv_agent))
        h = bridge.get_handle(n_agents, max_cell + 1, M, self.params, self.device)
        bridge.upload(h, dense, cell_type, cell_indid)
        replay = self.mode == "replay"
        if replay:
            if self.tape is None:
                raise RuntimeError("replay mode needs the reference's draw tape")
            t = self.tape
            h.set_tape(t.sel_key, t.sel_off, t.sel_ids, t.pair_k1, t.pair_k2, t.pair_u, t.expo_agent, t.expo_row)
        dp = self.dyn_params
        d = h.make_day(capi.MODE_REPLAY if replay else capi.MODE_PHILOX, int(day),
                       float(dp.masks_hygiene_distancing_multiplier), bool(dp.testing_enabled),
                       bool(dp.quarantine_enabled), seed=self.seed)
        h.simulate_contact_network(d, iv_agent, iv_cell, iv_start, iv_end)
        if transmit:
            h.simulate_direct_contacts(d)
        self.counters = h.counters()
        if replay and self.counters.tape_miss:
            raise RuntimeError("replay tape does not match this cell (%d misses)" % self.counters.tape_miss)
        rows, dur, _ = h.get_contacts()
        # the dict the reference hands on: {(a1, a2): [(start, end), ...]} in creation order, the overlap list in the
        # order of its pair loop (a1's intervals outer, a2's inner, contactnetwork.py:160-196).  The lists are a
        # re-formatting of the caller's own interval list; each is checked against what the device computed for the
        # row (first overlap's start, last overlap's end, summed duration).
        by_agent = {}
        for a, s0, e0 in lst:
            by_agent.setdefault(int(a), []).append((int(s0), int(e0)))
        for (sct, a1, a2, st, en), du in zip(rows.tolist(), dur.tolist()):
            ov = [(max(s1, s2), min(e1, e2)) for (s1, e1) in by_agent[a1] for (s2, e2) in by_agent[a2]
                  if max(s1, s2) <= min(e1, e2)]
            if not ov or ov[0][0] != st or ov[-1][1] != en or sum(e - s + 1 for s, e in ov) != du:
                raise RuntimeError("device contact row (%d, %d) disagrees with the cell's interval list" % (a1, a2))
            contacts[(a1, a2)] = ov
        if transmit:
            exposed = [int(a) for a in h.get_exposed()]
            if exposed:
                bridge.scatter(h.get_state(), exposed, self.agents_epi, vu)
        return contacts, exposed

    def generate_contact_network(self, cellid):
        """contactnetwork.py:134-315 -> (agents_directcontacts, cell_type)"""
        contacts, _ = self._run_cell(cellid, getattr(self, "_day", 0), False)
        return contacts, self.cells_type[cellid]

    def simulate_contact_network_by_cellid(self, cellid, day):
        """contactnetwork.py:127-132 -> (updated_agents_ids, agents_directcontacts, cell_type)"""
        self._day = int(day)
        contacts, exposed = self._run_cell(cellid, day, True)
        return exposed, contacts, self.cells_type[cellid]

    def simulate_contact_network(self, day, weekday):
        """contactnetwork.py:65-124 -> (process_index, updated_agents_ids, agents_partial, vars_util)"""
        vu = self.vars_util
        self._day = int(day)
        start = time.time()
        if self.process_index != -1:
            out_rows = []
        else:
            out_rows = vu.directcontacts_by_simcelltype_by_day
        # flatten {cellid: [(agent, start, end), ...]} in dict order = emission order (itinerary.py:1446-1452)
        iv_cell, iv_agent, iv_start, iv_end = bridge.flatten_intervals(vu.cells_agents_timesteps)
        updated_agents_ids = []
        M = len(iv_cell)
        if M > 0:
            n_agents = int(self.n_locals + self.n_tourists)
            max_cell = max(int(iv_cell.max()), max((int(c) for c in self.cells_type.keys()), default=0))
            cell_type, cell_indid = self._cell_tables(max_cell)
            active = np.unique(iv_agent)
            dense = bridge.flatten(n_agents, self.agents_static, self.agents_epi, vu, active, n_locals=self.n_locals)
            h = bridge.get_handle(n_agents, max_cell + 1, M, self.params, self.device)
            bridge.upload(h, dense, cell_type, cell_indid)
            replay = self.mode == "replay"
            if replay:
                if self.tape is None:
                    raise RuntimeError("replay mode needs the reference's draw tape")
                t = self.tape
                h.set_tape(t.sel_key, t.sel_off, t.sel_ids, t.pair_k1, t.pair_k2, t.pair_u, t.expo_agent, t.expo_row)
            dp = self.dyn_params
            d = h.make_day(capi.MODE_REPLAY if replay else capi.MODE_PHILOX, int(day),
                           float(dp.masks_hygiene_distancing_multiplier), bool(dp.testing_enabled),
                           bool(dp.quarantine_enabled), seed=self.seed)
            h.simulate_contact_network(d, iv_agent, iv_cell, iv_start, iv_end)
            h.simulate_direct_contacts(d)
            self.counters = h.counters()              # raises on device-side errors (capacity, oversized cell)
            if replay and self.counters.tape_miss:
                raise RuntimeError("replay tape does not match this day (%d misses)" % self.counters.tape_miss)
            rows, _, _ = h.get_contacts()
            out_rows.extend(rows.tolist())            # [simcelltype, a1, a2, first start, last end] (:98-105)
            updated_agents_ids = [int(a) for a in h.get_exposed()]
            if updated_agents_ids:
                bridge.scatter(h.get_state(), updated_agents_ids, self.agents_epi, vu)
        self.contactnetwork_sum_time_taken += time.time() - start
        agents_partial = {a: self.agents_epi[a] for a in updated_agents_ids}
        if self.process_index != -1:
            vu.directcontacts_by_simcelltype_by_day = out_rows
        return self.process_index, updated_agents_ids, agents_partial, vu
