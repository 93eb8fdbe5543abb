"""Tensor-native day loop of the whole state step (the hot-path part of sim.py:1392-2125) for the throughput
configurations (BASELINE.json configs 2-4): every stage runs on the device, in the reference's order --

    SEIR advance + vaccination dose day (itinerary.py:399, 2489-2527)          -> K4 (+K7 dose)
    intervention bookkeeping: test results -> contact_tracing_agent_ids, expiry
      of test / quarantine / hospitalisation records (itinerary.py:2283-2528)  -> K6
    contact network + transmission (contactnetwork.py:65, epidemiology.py:88) -> K0 (intervention mask) K1 K2 K3
    contact tracing over the last days_back + 1 days (sim.py:1929-1968)        -> K5
    vaccination scheduling (sim.py:2014, epidemiology.py:911-976)              -> K7
    refresh_dynamic_parameters at the END of the day (sim.py:2076; occupancy
      histogram from the device): the switches govern the next day            -> mhd, testing / quarantine / tracing

The itinerary stage itself is out of scope: the day's interval stream comes from the recorded bank (synth.py), and the
K0 intervention mask stands in for the stage's quarantine / hospital / deceased handling.  torch is plumbing here
(device buffers, the 3-day ring of contact rows); all arithmetic is in libmtd_b200.so.
"""
from __future__ import annotations

import collections

import numpy as np
import torch

from . import capi, params as prm, util
from .dynamicparams import DynamicParams, seir_states_from_occupancy


class EpidemicRun:
    """rank / world / group: one process per GPU, cells sharded over the ranks by the population's owner map, per-agent
    state replicated (the reference's stateful-actor strategy, actor_dist.py:432-667).  Two exchange steps per day: the
    exposure records of K3 (NVLink peer stores, engine.setup_peer_exchange) and -- on days with tracing work -- the
    primary-contact flags of K5's scan (one MAX all-reduce of N bytes; sim.py:2034-2063 syncs the traced agents between
    partitions at the same point).  K4, K6 and K7 run on every rank on the replicated state with Philox draws keyed by
    (day, agent): identical everywhere, no exchange.  The N-GPU state equals the 1-GPU state bit for bit."""

    def __init__(self, pop, epidemiologyparams, device=0, seed=0, max_rows=0, seed_initial_cases=True, rank=0, world=1,
                 group=None):
        self.pop, self.ep, self.seed = pop, epidemiologyparams, seed
        self.rank, self.world, self.group = rank, world, group
        self.dev = torch.device("cuda", device)
        self.P = prm.build_params(None, epidemiologyparams)
        self.streams = [pop.stream(k, rank, world) for k in range(pop.n_days)]
        max_m = max(len(s["iv_agent"]) for s in self.streams) + 16
        self.stream = torch.cuda.Stream(self.dev)
        self.delta_cap = max(4096, pop.n_agents // 16)
        self.h = capi.Handle(pop.n_agents, pop.n_cells, max_m, device=device, stream=self.stream.cuda_stream,
                             max_rows=max_rows, max_exposures=self.delta_cap)
        h = self.h
        h.set_params(self.P)
        self.exchange, self.gather = "none", None
        if world > 1:
            from . import engine
            self.exchange = engine.setup_peer_exchange(h, rank, world, self.delta_cap, group)
            if self.exchange == "nccl":
                with torch.cuda.device(self.dev):
                    self.gather = torch.zeros((world, self.delta_cap + 1, capi.DELTA_WORDS), dtype=torch.int32, device=self.dev)
            self.prim_flags = torch.zeros(pop.n_agents, dtype=torch.uint8, device=self.dev)
        D = capi.MEM_DEVICE
        with torch.cuda.stream(self.stream):
            h.set_cells(pop.cell_type, pop.cell_indid, mem=D)
            h.set_static(pop.age_bracket, pop.epi_age_bracket, pop.soc_rate, pop.res_cellid, mem=D)
            if seed_initial_cases:      # schedules come from K9 below, not from the population's placeholder entries
                h.set_state(seir=pop.seir, inf_type=pop.inf_type, inf_sev=pop.inf_sev, sched=torch.zeros_like(pop.sched),
                            sched_cnt=torch.zeros_like(pop.sched_cnt), mem=D)
            else:
                h.set_state(seir=pop.seir, inf_type=pop.inf_type, inf_sev=pop.inf_sev, sched=pop.sched, sched_cnt=pop.sched_cnt, mem=D)
        h.sync()
        self.n_seeded = 0
        if seed_initial_cases:
            # itinerary.py:373-395: on day 1 every agent that starts Infectious was exposed during the last
            # pre_exposure_days days; its whole trajectory is sampled from that day and its states reset (K9)
            ids = torch.nonzero(pop.seir == 3).flatten().to(torch.int32).cpu().numpy()
            d1 = h.make_day(capi.MODE_PHILOX, 1, 0.0, False, False, seed=seed)
            h.seed_exposures(d1, ids, pre_exposure_days=int(epidemiologyparams.get("pre_exposure_days", 7)))
            self.n_seeded = len(ids)
        h.set_intervention_mask(True)
        # vaccination: ids that are not part of the population are never candidates
        h.enable_vaccination(pop.n_agents)
        present = (pop.seir != 0)
        last = torch.where(present, prm.NONE, 0).to(torch.int32).cpu().numpy()
        h.set_vaccination(last, np.full((pop.n_agents, 2), prm.NONE, np.int32))
        self.n_present = int(present.sum())
        self.epi_present = present.to(torch.uint8).contiguous()
        # residences for the secondary contacts of K5: members of every residence cell (CSR over cell ids)
        rc = pop.res_cellid.to(torch.int64)
        ok = present & (rc >= 0)
        ids = torch.nonzero(ok).flatten()
        order = torch.argsort(rc[ids], stable=True)
        members = ids[order].to(torch.int32).contiguous()
        cnt = torch.bincount(rc[ids], minlength=pop.n_cells)
        off = torch.zeros(pop.n_cells + 1, dtype=torch.int64, device=self.dev)
        off[1:] = torch.cumsum(cnt, 0)
        with torch.cuda.stream(self.stream):
            h.set_residences(off, cnt.to(torch.int32).contiguous(), members, mem=D)
        h.sync()
        self.dyn = DynamicParams(self.n_present, 0, 0, epidemiologyparams)
        self.days_back = int(self.P.contact_tracing_days_back)
        self.ring = collections.deque(maxlen=self.days_back + 1)          # (day, rows3 device tensor)
        self.rows5 = torch.empty((max_rows or 4 * max_m + 1024, 5), dtype=torch.int32, device=self.dev)
        self.traced = torch.empty(pop.n_agents, dtype=torch.int32, device=self.dev)
        self.scheduled = collections.deque([0, 0], maxlen=2)              # doses land the day after scheduling
        self.n_vaccinated = 0
        self._rs = np.random.RandomState(7919 + int(seed))               # host-side scalar draws (doses per day)
        self.log = []

    def day(self, day):
        h, dp = self.h, self.dyn
        with torch.cuda.stream(self.stream):
            if day == 1 or not self.log:     # sim.py:1405-1407: the first day starts with a refresh of its own
                dp.refresh_dynamic_parameters(day, seir_states=seir_states_from_occupancy(h.occupancy()))
            d = h.make_day(capi.MODE_PHILOX, day, float(dp.masks_hygiene_distancing_multiplier), dp.testing_enabled,
                           dp.quarantine_enabled, seed=self.seed)
            # the itinerary stage advances an agent's state first (itinerary.py:399,1587) and then does the intervention
            # bookkeeping on the new state (sample_intervention_activities, :1282 -> :2283-2528)
            h.advance_seir(day, None)
            n_traced = h.process_test_results_device(d, self.traced, self.epi_present)
            self.n_vaccinated += self.scheduled[0]                        # scheduled two calls ago -> dosed by now
            s = self.streams[day % self.pop.n_days]
            h.simulate_contact_network(d, s["iv_agent"], s["iv_cell"], s["iv_start"], s["iv_end"], iv_order=s.get("iv_order"),
                                       mem=capi.MEM_DEVICE)
            self._transmission(d)
            c = h.counters()
            n_rows = h.get_contacts_device(self.rows5)
            self.ring.appendleft((day, self.rows5[:n_rows, :3].contiguous()))
            st = None
            if dp.contact_tracing_enabled and n_traced > 0:
                slices = [r for (_, r) in self.ring]
                off = np.concatenate([[0], np.cumsum([len(r) for r in slices])]).astype(np.int64)
                rows3 = torch.cat(slices) if len(slices) > 1 else slices[0]
                if self.world == 1:
                    h.contact_tracing(d, self.traced[:n_traced], off, rows3, epi_present=self.epi_present, mem=capi.MEM_DEVICE)
                else:       # scan this rank's rows, merge the primary-contact flags, schedule on the replicated records
                    h.contact_tracing_scan(d, self.traced[:n_traced], off, rows3, self.prim_flags, epi_present=self.epi_present)
                    torch.distributed.all_reduce(self.prim_flags, op=torch.distributed.ReduceOp.MAX, group=self.group)
                    h.contact_tracing_finish(self.prim_flags)
                st = h.trace_stats()
            n_sched = 0
            if dp.vaccination_propensity > 0:
                n_sched = self._schedule_vaccinations(d)
            self.scheduled.append(n_sched)
            occ = h.occupancy()
        pairs, n_rows_day = int(c.pairs_evaluated), int(c.n_rows)
        if self.world > 1:          # the log line is global; everything else in it is replicated state
            t = torch.tensor([pairs, n_rows_day], dtype=torch.int64, device=self.dev)
            torch.distributed.all_reduce(t, group=self.group)
            pairs, n_rows_day = int(t[0]), int(t[1])
        row = dict(day=day, occupancy=[int(x) for x in occ], pairs=pairs, rows=n_rows_day,
                   exposed=int(c.n_exposed), traced=n_traced, primary=int(st.n_primary) if st else 0,
                   secondary=int(st.n_secondary) if st else 0, vaccinations=n_sched, sched_overflow=int(c.sched_overflow),
                   mhd=float(dp.masks_hygiene_distancing_multiplier), testing=bool(dp.testing_enabled),
                   quarantine=bool(dp.quarantine_enabled), tracing=bool(dp.contact_tracing_enabled))
        self.log.append(row)
        # end of the day (sim.py:2076): the switches computed here, with today's day number and the occupancy after
        # today, govern TOMORROW
        dp.refresh_dynamic_parameters(day, seir_states=seir_states_from_occupancy(occ))
        return row

    def _transmission(self, d):
        """K3 of the day + the exchange and application of the exposure records (engine.DayEngine._transmission)"""
        h = self.h
        if self.world == 1:
            h.simulate_direct_contacts(d)
        elif self.exchange == "p2p":
            h.simulate_direct_contacts(d, apply=capi.APPLY_EXCHANGE)
        else:
            from . import engine
            h.simulate_direct_contacts(d, delta_buf=self.gather[self.rank], cap_records=self.delta_cap, apply=False)
            engine.exchange_deltas(self.gather, self.rank, self.group)
            h.apply_deltas(self.gather, self.world, self.delta_cap)

    def _schedule_vaccinations(self, d):
        """the dyn_params bookkeeping of epidemiology.py:917-945 on host scalars; the sample on the device (K7)"""
        dp = self.dyn
        n_locals = self.n_present
        not_vacc = n_locals - self.n_vaccinated
        if dp.vaccination_propensity != dp.last_vaccination_propensity:
            if dp.num_agents_to_be_vaccinated == 0:
                dp.num_agents_to_be_vaccinated = round(not_vacc * dp.vaccination_propensity) or 1
            else:
                dp.num_agents_to_be_vaccinated += round(not_vacc * (dp.vaccination_propensity - dp.last_vaccination_propensity)) or 1
            dp.last_vaccination_propensity = dp.vaccination_propensity
        mean, std = self.ep["vaccination_daily_parameters"]
        ratio = n_locals / 519562.0                                     # locals_ratio_to_full_pop (sim.py: Malta's population)
        mu, sigma = util.lognormal_params(mean, std)                   # util.sample_log_normal(mean, std, 1, True) on a stream of
        drawn = round(float(self._rs.lognormal(mu, sigma)))             #   this run's own: every rank draws the same number
        dp.num_vaccinations_today = min(round(drawn * ratio),
                                        dp.num_agents_to_be_vaccinated - self.n_vaccinated)
        if dp.num_vaccinations_today <= 0:
            return 0
        return self.h.schedule_vaccinations_count(d, int(dp.num_vaccinations_today))

    def close(self):
        self.h.close()
