"""Drop-in for the hot-path methods of the reference's `epidemiology.Epidemiology` (epidemiology.py:14-906): same
constructor; `contact_tracing(day, distributed=False, f=None)` (epidemiology.py:667-834) and
`contact_tracing_clean_up(day, f=None)` (:838-906) computed through libmtd_b200.so (K5).  The transmission step
(`simulate_direct_contacts`, :88-231) runs fused behind `ContactNetwork.simulate_contact_network` (K3) and can also be
called on its own with a dict of contacts, like the reference's; there is no CPU fallback for any of it.

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
        self.cn_tape = None            # replay tape of the contact-network stage (set by ContactNetwork)
        self._params = None

    @property
    def params(self):
        if self._params is None:
            self._params = prm.build_params(None, self.epidemiologyparams)
        return self._params

    def simulate_direct_contacts(self, agents_directcontacts, cellid, cell_type, day):
        """epidemiology.py:88-231 -> list of newly exposed agent ids, in the order of the contact dict.
        The pairs and their overlap lists go to the device as they are (mtd_set_contacts); K3 evaluates them like the
        rows of the fused day step.  (Inside ContactNetwork.simulate_contact_network the rows never leave HBM.)"""
        from . import bridge, capi
        if not agents_directcontacts:
            return []
        vu = self.vars_util
        pairs = list(agents_directcontacts.items())
        a1 = np.array([int(k[0]) for k, _ in pairs], np.int32)
        a2 = np.array([int(k[1]) for k, _ in pairs], np.int32)
        off = np.zeros(len(pairs) + 1, np.int64)
        se = []
        for i, (_, ov) in enumerate(pairs):
            for (s0, e0) in ov:
                se.append(int(s0)); se.append(int(e0))
            off[i + 1] = len(se) // 2
        se = np.array(se, np.uint8)
        n_agents = int(self.n_locals + self.n_tourists)
        cellid = int(cellid)
        ct = np.full(cellid + 1, -1, np.int8); ct[cellid] = int(cell_type)
        ci = np.full(cellid + 1, -1, np.int16)
        dense = bridge.flatten(n_agents, self.agents_static, self.agents_epi, vu, np.unique(np.concatenate([a1, a2])))
        h = bridge.get_handle(n_agents, cellid + 1, 16, self.params, self.device)
        bridge.upload(h, dense, ct, ci)
        replay = self.mode == "replay"
        if replay:
            if self.cn_tape is None:
                raise RuntimeError("replay mode needs the reference's draw tape")
            t = self.cn_tape
            h.set_tape(t.sel_key, t.sel_off, t.sel_ids, t.pair_k1, t.pair_k2, t.pair_u, t.expo_agent, t.expo_row)
        dp = self.dyn_params
        d = h.make_day(capi.MODE_REPLAY if replay else capi.MODE_PHILOX, int(day),
                       float(dp.masks_hygiene_distancing_multiplier), bool(dp.testing_enabled),
                       bool(dp.quarantine_enabled), seed=self.seed)
        h.set_contacts(d, cellid, a1, a2, off, se)
        h.simulate_direct_contacts(d)
        h.counters()                                   # raises on device-side errors
        exposed = [int(a) for a in h.get_exposed()]
        if exposed:
            bridge.scatter(h.get_state(), exposed, self.agents_epi, vu)
        return exposed

    # ------------------------------------------------------------------ trajectory of one exposure (K9)
    def simulate_seir_state_transition(self, current_day, exposed_agent, exposed_agent_id, incremental_days,
                                       overlapping_timesteps, agent_state_transition_by_day, agent_epi_age_bracket_index,
                                       agent_quarantine_days, asymptomatic_multiplier=1.0):
        """epidemiology.py:233-500 called on its own, as the itinerary stage does for the agents that start Infectious
        (itinerary.py:373-395: `incremental_days` < 1 -> reset_states).  -> (agent_state_transition_by_day, seir_state,
        infection_type, infection_severity, recovered); test and hospitalisation records are written into
        `exposed_agent`, the quarantine record is left to the caller's overwrite (itinerary.py:390), as in the
        reference.  Sampled on the device (mtd_seed_exposures); replay mode reads one EXPO_W row from `self.tape`.
        For many agents use `seed_initial_cases`, one launch for the whole batch."""
        from . import bridge, capi
        from .epidemiologyclasses import InfectionType, SEIRState, SEIRStateTransition, Severity
        a = int(exposed_agent_id)
        n_agents = max(int(self.n_locals + self.n_tourists), a + 1)
        st = bridge.DenseState(n_agents)
        st.epi_age_bracket[a] = int(agent_epi_age_bracket_index)
        bridge._rec(st.test_day, a, exposed_agent.get("test_day")); bridge._rec(st.test_result_day, a, exposed_agent.get("test_result_day"))
        bridge._rec(st.hosp, a, exposed_agent.get("hospitalisation_days"))
        h = bridge.get_handle(n_agents, 1, 16, self.params, self.device)
        bridge.upload(h, st, np.zeros(1, np.int8), np.full(1, -1, np.int16))
        replay = self.mode == "replay"
        if replay and self.tape is None:
            raise RuntimeError("replay mode needs the reference's draws of this call (one EXPO_W row)")
        dp = self.dyn_params
        d = h.make_day(capi.MODE_REPLAY if replay else capi.MODE_PHILOX, int(current_day), 0.0, bool(dp.testing_enabled),
                       bool(dp.quarantine_enabled), seed=self.seed)
        out = h.seed_exposures(d, [a], exposure_day=[int(incremental_days)], timesteps=np.asarray(overlapping_timesteps),
                               asym_mult=[float(asymptomatic_multiplier)],
                               draws=np.asarray(self.tape, np.float64).reshape(1, -1) if replay else None)[0]
        dev = h.get_state()
        if agent_state_transition_by_day is None:
            agent_state_transition_by_day = []
        for v in dev["sched"][a, :int(dev["sched_cnt"][a])]:
            day, trans, ts = bridge.unpack_sched(v)
            agent_state_transition_by_day.append([day, SEIRStateTransition(trans), ts])
        if dev["test_day"][a, 0] != prm.NONE and st.test_day[a, 0] == prm.NONE:
            exposed_agent["test_day"] = bridge._lst(dev["test_day"][a])
            exposed_agent["test_result_day"] = bridge._lst(dev["test_result_day"][a])
        if dev["hosp"][a, 0] != prm.NONE:
            exposed_agent["hospitalisation_days"] = bridge._lst(dev["hosp"][a])
        return (agent_state_transition_by_day, SEIRState(int(out[0])), InfectionType(int(out[1])), Severity(int(out[2])),
                bool(out[3]))

    def seed_initial_cases(self, simday, agent_ids, pre_exposure_days, potential_timesteps=None):
        """The loop of itinerary.py:373-395 over all agents that start Infectious, as ONE device batch: exposure day drawn
        from arange(simday - pre_exposure_days, simday), trajectory sampled, states reset, records written back into
        agents_epi / vars_util.  -> the ids, in the order given."""
        from . import bridge, capi
        ids = np.ascontiguousarray([int(a) for a in agent_ids], np.int32)
        if len(ids) == 0:
            return []
        n_agents = int(self.n_locals + self.n_tourists)
        dense = bridge.flatten(n_agents, self.agents_static, self.agents_epi, self.vars_util, ids)
        h = bridge.get_handle(n_agents, 1, 16, self.params, self.device)
        bridge.upload(h, dense, np.zeros(1, np.int8), np.full(1, -1, np.int16))
        dp = self.dyn_params
        d = h.make_day(capi.MODE_PHILOX, int(simday), 0.0, bool(dp.testing_enabled), bool(dp.quarantine_enabled), seed=self.seed)
        h.seed_exposures(d, ids, pre_exposure_days=int(pre_exposure_days),
                         timesteps=np.arange(144) if potential_timesteps is None else np.asarray(potential_timesteps))
        dev = h.get_state()
        quar = {int(a): self.agents_epi[int(a)].get("quarantine_days") for a in ids}
        bridge.scatter(dev, ids, self.agents_epi, self.vars_util)
        for a, q in quar.items():                       # agent_epi["quarantine_days"] = agent_quarantine_days (:390)
            self.agents_epi[a]["quarantine_days"] = q
        return [int(a) for a in ids]

    # ------------------------------------------------------------------ small host-side helpers of the class
    # The itinerary stage calls the three schedulers below once per agent on ITS side of the boundary (itinerary.py:2319,
    # 2330, 2383; SURVEY.md section 8b "Who calls it") with the legacy global RNG streams, so they are host methods on the
    # agent's dict that consume random.random() / np.random.lognormal exactly like the reference's.  The day step's own
    # uses of the same rules (exposures of K3, traced contacts of K5, positive results of K6) run on the device.
    def schedule_test(self, agent, agent_id, incremental_days, start_timestep, quarantine_type):
        """epidemiology.py:502-565 -> (scheduled, test_day, test_result_day)"""
        import random
        from . import util
        from .epidemiologyclasses import QuarantineType
        ep = self.epidemiologyparams
        test_day = result_day = None
        if not self.dyn_params.testing_enabled:
            return False, test_day, result_day
        test_day = agent["test_day"]
        if test_day is not None and len(test_day) > 0:                  # one pending test at a time (:519-521)
            return False, test_day, result_day
        if quarantine_type == QuarantineType.Symptomatic and not (random.random() < ep["testing_after_symptoms_probability"]):
            return False, test_day, result_day
        wait = {QuarantineType.PositiveContact: self.contact_tracing_positive_delay_days,
                QuarantineType.SecondaryContact: self.contact_tracing_secondary_delay_days}.get(quarantine_type, 0)
        wait += util.sample_log_normal(*ep["testing_days_distribution_parameters"][:2], 1, True)
        day_of_test = incremental_days + wait
        until_result = util.sample_log_normal(*ep["testing_results_days_distribution_parameters"][:2], 1, True)
        agent["test_day"] = [day_of_test, start_timestep]
        agent["test_result_day"] = [day_of_test + until_result, start_timestep]
        return True, agent["test_day"], agent["test_result_day"]

    def schedule_quarantine(self, agent_id, start_day, start_timestep, quarantine_type, end_day=None, quarantine_days=None,
                            agent=None):
        """epidemiology.py:568-625 -> (scheduled, quarantine_days)"""
        import random
        from .epidemiologyclasses import QuarantineType
        ep = self.epidemiologyparams
        if not self.dyn_params.quarantine_enabled:
            return False, quarantine_days
        u = random.random()
        if quarantine_type == QuarantineType.Symptomatic:
            comply = u < ep["quarantine_after_symptoms_probability"]
        else:
            comply = not (u < ep["ignore_quarantine_rules_probability"])
        if not comply:
            return False, quarantine_days
        record = agent if agent is not None else self.agents_epi[agent_id]
        if quarantine_days is None:
            quarantine_days = record["quarantine_days"]
        if end_day is None:
            d_pos, d_contact, d_secondary = ep["quarantine_duration_parameters"][:3]
            if quarantine_type in (QuarantineType.Symptomatic, QuarantineType.Positive):
                end_day = start_day + d_pos
            elif quarantine_type == QuarantineType.PositiveContact:
                start_day += self.contact_tracing_positive_delay_days
                end_day = start_day + d_contact
            elif quarantine_type == QuarantineType.SecondaryContact:
                start_day += self.contact_tracing_secondary_delay_days
                end_day = start_day + d_secondary
        if quarantine_days is not None and len(quarantine_days) > 0 and start_day >= quarantine_days[0]:
            return False, quarantine_days                               # never moved to the same or a later start (:606-611)
        quarantine_days = [start_day, start_timestep, end_day]
        record["quarantine_days"] = quarantine_days
        return True, quarantine_days

    def schedule_hospitalisation(self, agent, agent_id, hospitalisation_days):
        """epidemiology.py:647-660: an existing window keeps its start day and takes the new end"""
        cur = agent["hospitalisation_days"]
        if cur is not None and len(cur) > 0:
            agent["hospitalisation_days"] = [cur[0], hospitalisation_days[1], hospitalisation_days[2]]
        else:
            agent["hospitalisation_days"] = hospitalisation_days
        # (returns None, like the reference's)

    def update_quarantine(self, agent, new_start_day, new_start_ts, new_end_day, new_end_ts):
        """epidemiology.py:627-630: overwrite the agent's quarantine record [start day, timestep, end day]"""
        agent["quarantine_days"] = [new_start_day, new_start_ts, new_end_day]

    def convert_celltype_to_base_infection_prob(self, cellid, rescellid, celltype=None):
        """epidemiology.py:998-1014 (the table K3 applies per row; quirk kept: a workplace counts as "workplace" only
        when it is the exposed agent's RESIDENCE cell)"""
        from . import util
        from .epidemiologyclasses import SimCellType
        P = self.params
        sct = util.convert_celltype_to_simcelltype(cellid, celltype=celltype)
        if sct == SimCellType.Residence:
            return P.p_household
        if sct == SimCellType.Workplace:
            return P.p_workplace if cellid == rescellid else P.p_community
        if sct == SimCellType.School:
            return P.p_school
        return P.p_community

    def convert_simcelltype_to_contact_tracing_success_prob(self, simcelltype):
        """epidemiology.py:978-996 (the table K5 applies per scanned row)"""
        i = int(simcelltype)
        return self.params.tracing_success_prob[i] if 0 <= i <= 3 else self.params.tracing_success_prob[3]

    # ------------------------------------------------------------------ vaccination (K7)
    def schedule_vaccinations(self, day):
        """epidemiology.py:911-976 -> ids of the agents scheduled for day + 1.  The bookkeeping on dyn_params
        (num_agents_to_be_vaccinated, num_vaccinations_today; :917-945) is host arithmetic on scalars, as in the
        reference (one log-normal draw from the legacy stream); the draw of the agents and of their timesteps runs on
        the device (mtd_schedule_vaccinations)."""
        from . import bridge, capi, util
        dp, vu = self.dyn_params, self.vars_util
        updated = []
        if dp.vaccination_propensity > 0:
            doses = vu.agents_vaccination_doses
            num_not_vaccinated = self.n_locals - len(doses)
            num_already_vaccinated = self.n_locals - num_not_vaccinated
            if dp.vaccination_propensity != dp.last_vaccination_propensity:
                if dp.num_agents_to_be_vaccinated == 0:
                    dp.num_agents_to_be_vaccinated = round(num_not_vaccinated * dp.vaccination_propensity) or 1
                else:
                    change = dp.vaccination_propensity - dp.last_vaccination_propensity
                    dp.num_agents_to_be_vaccinated += round(num_not_vaccinated * change) or 1
                dp.last_vaccination_propensity = dp.vaccination_propensity
            mean, std = self.epidemiologyparams["vaccination_daily_parameters"]
            dp.num_vaccinations_today = round(util.sample_log_normal(mean, std, 1, True) * self.locals_ratio_to_full_pop)
            remaining = dp.num_agents_to_be_vaccinated - num_already_vaccinated
            if dp.num_vaccinations_today > remaining:
                dp.num_vaccinations_today = remaining
            if dp.num_vaccinations_today > 0:
                n_agents = int(self.n_locals + self.n_tourists)
                h = bridge.get_handle(n_agents, 1, 16, self.params, self.device)
                h.enable_vaccination(int(self.n_locals))
                last = np.full(n_agents, prm.NONE, np.int32)
                for a, d in doses.items():
                    if int(a) < n_agents and len(d) > 0:
                        last[int(a)] = int(d[-1])
                h.set_vaccination(last, np.full((n_agents, 2), prm.NONE, np.int32))
                replay = self.mode == "replay"
                d = h.make_day(capi.MODE_REPLAY if replay else capi.MODE_PHILOX, int(day), 0.0, False, False, seed=self.seed)
                if replay:
                    if self.tape is None:
                        raise RuntimeError("replay mode needs the reference's vaccination sample (ids, timesteps)")
                    ids, ts = self.tape
                    h.schedule_vaccinations(d, len(ids), ids, ts)
                    order = [int(a) for a in ids]
                else:
                    order = [int(a) for a in h.schedule_vaccinations(d, int(dp.num_vaccinations_today))]
                _, rec = h.get_vaccination()
                for a in order:
                    e = self.agents_epi[a]
                    lst = e.get("vaccination_days") or []
                    lst.append([int(rec[a, 0]), int(rec[a, 1])])           # [day + 1, sampled_timestep] (:965)
                    e["vaccination_days"] = lst
                    updated.append(a)
        return updated

    # ------------------------------------------------------------------ contact tracing (K5)
    def _residences(self):
        """cells_households / cells_institutions / cells_accommodation -> CSR over cell ids: members = residents
        (resident_uids / member_uids) followed by staff (staff_uids), as epidemiology.py:752-788 concatenates them"""
        cells = {}
        for src, rk, sk in ((self.cells_households, "resident_uids", "staff_uids"),
                            (self.cells_institutions, "resident_uids", "staff_uids"),
                            (self.cells_accommodation, "member_uids", None)):
            for cid, cell in (src or {}).items():
                if int(cid) in cells:                       # households before institutions before accommodation
                    continue
                place = cell["place"]
                res = place.get(rk) if hasattr(place, "get") else place[rk]
                stf = (place.get(sk) if hasattr(place, "get") else place[sk]) if sk else None
                cells[int(cid)] = ([int(x) for x in (res if res is not None else [])],
                                   [int(x) for x in (stf if stf is not None else [])])
        n = (max(cells) + 1) if cells else 1
        off = np.zeros(n + 1, np.int64); nres = np.zeros(n, np.int32); mem = []
        for c in range(n):
            r, s = cells.get(c, ([], []))
            mem.extend(r); mem.extend(s); nres[c] = len(r); off[c + 1] = len(mem)
        return off, nres, np.array(mem, np.int32)

    def contact_tracing(self, day, distributed=False, f=None):
        """epidemiology.py:667-834 -> (process_index, updated_agent_ids, agents_epi, vars_util)"""
        from . import bridge, capi
        vu = self.vars_util
        updated = []
        traced = sorted(int(x) for x in vu.contact_tracing_agent_ids)
        if self.dyn_params.contact_tracing_enabled and len(traced) > 0:
            if getattr(self.dyn_params, "statistics", None) is not None:
                self.dyn_params.statistics.new_contacttraced = len(traced)
            rows = vu.directcontacts_by_simcelltype_by_day
            markers = vu.directcontacts_by_simcelltype_by_day_start_marker
            slices, off = [], [0]
            for back in range(self.contact_tracing_days_back + 1):          # :677-703, start markers as the reference
                d = day - back
                sl = []
                if d in markers:
                    lo, hi = markers[d], len(rows)
                    if d + 1 in markers:
                        hi = markers[d + 1] - 1                               # exclusive bound, as written (quirk B8)
                    sl = rows[lo:hi]
                slices.append(np.array([[int(r[0]), int(r[1]), int(r[2])] for r in sl], np.int32).reshape(-1, 3))
                off.append(off[-1] + len(slices[-1]))
            rows3 = np.ascontiguousarray(np.concatenate(slices)) if off[-1] else np.zeros((0, 3), np.int32)
            n_agents = int(self.n_locals + self.n_tourists)
            st = bridge.DenseState(n_agents)
            present = np.zeros(n_agents, np.uint8)
            keys = self.agents_epi.keys() if hasattr(self.agents_epi, "keys") else range(len(self.agents_epi))
            for a in keys:
                e = self.agents_epi[a]
                a = int(a)
                if e is None or len(e) == 0:
                    continue
                present[a] = 1
                bridge._rec(st.quarantine, a, e.get("quarantine_days")); bridge._rec(st.test_day, a, e.get("test_day"))
                bridge._rec(st.test_result_day, a, e.get("test_result_day"))
                try:
                    st.res_cellid[a] = int(self.agents_static.get(a, "res_cellid"))
                except Exception:
                    if a < self.n_locals:                                     # departed tourists are skipped (:744-748)
                        raise ValueError("Agent {0} not found during contact tracing".format(str(a)))
            off_r, nres, mem = self._residences()
            h = bridge.get_handle(n_agents, max(len(nres), 1), 16, self.params, self.device)
            h.set_static(st.age_bracket, st.epi_age_bracket, st.soc_rate, st.res_cellid)
            h.set_state(quarantine=st.quarantine, test_day=st.test_day, test_result_day=st.test_result_day)
            h.set_residences(off_r, nres, mem)
            replay = self.mode == "replay"
            if replay and self.tape is None:
                raise RuntimeError("replay mode needs the reference's contact-tracing tape")
            d = h.make_day(capi.MODE_REPLAY if replay else capi.MODE_PHILOX, int(day), 0.0,
                           bool(self.dyn_params.testing_enabled), bool(self.dyn_params.quarantine_enabled), seed=self.seed)
            h.contact_tracing(d, np.array(traced, np.int32), np.array(off, np.int64), rows3, epi_present=present,
                              tape=self.tape if replay else None)
            prim, sec, stats = h.get_traced()
            if replay and stats.tape_miss:
                raise RuntimeError("contact-tracing tape does not match this call (%d misses)" % stats.tape_miss)
            if getattr(self.dyn_params, "statistics", None) is not None and len(slices) > 0:
                self.dyn_params.statistics.num_direct_contacts = len(slices[-1])
            dev = h.get_state()
            prim_ids = [int(a) for a in np.nonzero(prim)[0]]
            for a in prim_ids:                                                # only primaries' records change (quirk B9)
                if not present[a]:
                    continue
                e = self.agents_epi[a]
                e["quarantine_days"] = bridge._lst(dev["quarantine"][a])
                e["test_day"] = bridge._lst(dev["test_day"][a]); e["test_result_day"] = bridge._lst(dev["test_result_day"][a])
            updated = prim_ids + [int(a) for a in np.nonzero(sec)[0]]
        if not distributed:
            self.contact_tracing_clean_up(day, f)
        return self.process_index, updated, self.agents_epi, vu

    def contact_tracing_clean_up(self, day, f=None):
        """epidemiology.py:838-906: drop the oldest retained day of contact rows and re-base the start markers
        (host-side list bookkeeping on the reference's own containers)"""
        vu = self.vars_util
        markers = vu.directcontacts_by_simcelltype_by_day_start_marker
        day_to_clean = day - self.contact_tracing_days_back
        day_end_idx = 0
        if day_to_clean + 1 in markers:
            day_end_idx = markers[day_to_clean + 1] - 1
        if day_end_idx > 0:
            vu.directcontacts_by_simcelltype_by_day = vu.directcontacts_by_simcelltype_by_day[day_end_idx + 1:]
            if day_to_clean in markers:
                del markers[day_to_clean]
            keys = list(markers.keys())
            rebased = {}
            for i in range(len(keys), 0, -1):                                 # as written at :886-900
                cur = keys[i - 1]
                prev = keys[i - 2] if i - 2 >= 0 else None
                rebased[cur] = 0 if prev is None else markers[cur] - markers[prev]
            vu.directcontacts_by_simcelltype_by_day_start_marker = rebased
