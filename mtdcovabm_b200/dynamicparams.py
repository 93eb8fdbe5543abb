"""Day-to-day switches of the hot path, refreshed from the DEVICE occupancy counts.

Mirror of the reference's `DynamicParams` (dynamicparams.py:6-172) and of the rates of `DynamicStatistics.refresh_rates`
(dynamicstatistics.py:59-142) for the fields the day step consumes -- `masks_hygiene_distancing_multiplier`,
`testing_enabled`, `quarantine_enabled`, `contact_tracing_enabled`, `vaccination_propensity` -- plus the lockdown and
travel-restriction switches the itinerary stage reads.  Same constructor and attribute names, so an instance can be
handed to `ContactNetwork` / `Epidemiology` as `dyn_params`.  The six SEIR counts come from `mtd_occupancy` (K6, a
histogram over the device-resident state: no device-to-host copy of the per-agent state), not from a Python loop over
`vars_util.agents_seir_state` (dynamicstatistics.py:169-193).
"""
from __future__ import annotations


class DynamicStatistics:
    """the rates of dynamicstatistics.py:59-142 (the CSV/log columns of the reference are out of scope)"""

    def __init__(self, n_locals, n_tourists, n_tourists_initial):
        self.n_locals, self.n_tourists, self.n_tourists_initial = n_locals, n_tourists, n_tourists_initial
        self.total_deceased = self.total_exposed = self.total_susceptible = self.total_infectious = self.total_recovered = 0
        self.new_deaths = self.new_exposed = self.new_susceptible = self.new_infectious = self.new_recovered = 0
        self.total_active_tourists = self.total_arriving_tourists = self.total_departing_tourists = 0
        self.total_arriving_nextday_tourists = 0
        self.total_active_population = n_locals
        self.infectious_rate = self.recovery_rate = self.mortality_rate = 0
        self.num_direct_contacts = 0
        self.avg_direct_contacts_per_agent = 0
        self.new_contacttraced = 0

    def refresh_rates(self, day, act_tourists, arr_tourists, arr_nextday_tourists, dep_tourists, seir_states):
        """seir_states = (deceased, exposed, susceptible, infectious, recovered), the reference's tuple order
        (dynamicstatistics.py:75)"""
        n_dec, n_exp, n_sus, n_inf, n_rec = (int(x) for x in seir_states)
        prev = (self.total_deceased, self.total_exposed, self.total_susceptible, self.total_infectious, self.total_recovered)
        new = [n if day == 1 else n - p for n, p in zip((n_dec, n_exp, n_sus, n_inf, n_rec), prev)]
        self.new_deaths, self.new_exposed, self.new_susceptible, self.new_infectious, self.new_recovered = new
        self.total_deceased, self.total_exposed, self.total_susceptible = n_dec, n_exp, n_sus
        self.total_infectious, self.total_recovered = n_inf, n_rec
        self.total_arriving_tourists, self.total_arriving_nextday_tourists = arr_tourists, arr_nextday_tourists
        self.total_departing_tourists, self.total_active_tourists = dep_tourists, act_tourists
        self.total_active_population = self.n_locals + act_tourists - n_dec
        if self.num_direct_contacts > 0:
            self.avg_direct_contacts_per_agent = round(self.num_direct_contacts / self.total_active_population, 2)
        self.infectious_rate = n_inf / self.total_active_population
        self.recovery_rate = n_rec / self.total_active_population
        self.mortality_rate = n_dec / self.total_active_population


def seir_states_from_occupancy(occ):
    """mtd_occupancy's counts indexed by SEIRState value (0 Undefined .. 5 Deceased) -> the reference's tuple
    (deceased, exposed, susceptible, infectious, recovered)"""
    return int(occ[5]), int(occ[2]), int(occ[1]), int(occ[3]), int(occ[4])


class DynamicParams:
    def __init__(self, n_locals, n_tourists, n_tourists_initial, epidemiologyparams):
        ep = epidemiologyparams
        self.n_locals, self.n_tourists, self.n_tourists_initial = n_locals, n_tourists, n_tourists_initial
        for group in ("masks_hygiene_distancing", "vaccination", "intervention", "travel_restrictions", "lockdown"):
            setattr(self, group + "_day_thresholds", ep[group + "_day_thresholds"])
            setattr(self, group + "_infectiousrate_thresholds", ep[group + "_infectiousrate_thresholds"])
        self.quarantine_enabled = self.testing_enabled = self.contact_tracing_enabled = False
        self.workplaces_lockdown = self.schools_lockdown = self.entertainment_lockdown = False
        self.activities_lockdown = self.airport_lockdown = False
        self.masks_hygiene_distancing_multiplier = 0
        self.travel_restrictions_multiplier = 1
        self.vaccination_propensity = 0
        self.last_vaccination_propensity = 0
        self.num_agents_to_be_vaccinated = 0
        self.num_vaccinations_today = 0
        self.statistics = DynamicStatistics(n_locals, n_tourists, n_tourists_initial)

    @staticmethod
    def get_value_by_rate_in_threshold(rate, params, none_val=0):
        """dynamicparams.py:151-169: value of the last threshold row [threshold, value] that `rate` has reached"""
        for i, (lo, val) in enumerate(params):
            if rate >= lo and (i + 1 == len(params) or rate < params[i + 1][0]):
                return val
        return none_val

    def _level(self, group, day):
        """(x, thresholds): the infectious rate against the rate thresholds when those are given, else the day against
        the day thresholds (dynamicparams.py:65-72)"""
        by_rate = getattr(self, group + "_infectiousrate_thresholds")
        if len(by_rate) == 0:
            return day, getattr(self, group + "_day_thresholds")
        return self.statistics.infectious_rate, by_rate

    def refresh_dynamic_parameters(self, day, num_act_tourists=0, num_arr_tourists=0, num_arr_nextday_tourists=0,
                                   num_dep_tourists=0, vars_util=None, force_infectious_rate_refresh=True,
                                   seir_states=None, handle=None):
        """dynamicparams.py:59-134.  `seir_states`: the reference's 5-tuple, or `handle`: a capi.Handle whose device
        state is counted by K6 (mtd_occupancy)."""
        if seir_states is None:
            if handle is None:
                raise RuntimeError("refresh_dynamic_parameters needs seir_states or a device handle (no CPU count loop)")
            seir_states = seir_states_from_occupancy(handle.occupancy())
        self.statistics.refresh_rates(day, num_act_tourists, num_arr_tourists, num_arr_nextday_tourists,
                                      num_dep_tourists, seir_states)
        x, th = self._level("masks_hygiene_distancing", day)
        self.masks_hygiene_distancing_multiplier = self.get_value_by_rate_in_threshold(x, th)
        x, th = self._level("vaccination", day)
        self.vaccination_propensity = self.get_value_by_rate_in_threshold(x, th)
        x, th = self._level("intervention", day)
        self.quarantine_enabled, self.testing_enabled, self.contact_tracing_enabled = (x >= th[0], x >= th[1], x >= th[2])
        x, th = self._level("travel_restrictions", day)
        self.travel_restrictions_multiplier = 1 - self.get_value_by_rate_in_threshold(x, th)
        x, th = self._level("lockdown", day)
        (self.workplaces_lockdown, self.schools_lockdown, self.entertainment_lockdown, self.activities_lockdown,
         self.airport_lockdown) = (x >= th[0], x >= th[1], x >= th[2], x >= th[3], x >= th[4])
