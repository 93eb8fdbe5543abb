"""Drop-in for the reference's `seirstateutil` module (seirstateutil.py:1-136).

`update_agent_state` keeps the reference's scalar signature (it is called once per agent from the itinerary stage,
itinerary.py:399,664,1587) and `update_agents_state` is the batched form the day engine uses; both run K4
(`mtd_advance_seir`) on the device -- the transition table of seirstateutil.py:105-136 lives only in the CUDA kernel.
`initialize_agent_states` (seirstateutil.py:5-32) is one-off host-side initialisation driven by the legacy global
RNGs, exactly like the reference's.
"""
from __future__ import annotations

import random

import numpy as np

from . import bridge, capi, params as prm
from .epidemiologyclasses import InfectionType, SEIRState, SEIRStateTransition, Severity

_scalar = {}


def _scalar_handle(device=0):
    h = _scalar.get(device)
    if h is None:
        h = capi.Handle(1, 1, 16, device=device)
        h.set_params(prm.build_params())
        _scalar[device] = h
    return h


def initialize_agent_states(n, initial_seir_state_distribution, agents_seir_state):
    """seirstateutil.py:5-32: round(n * fraction) agents per state drawn without replacement from the still-Undefined
    ones (np.random.choice), leftovers get random.randint(1, 4)"""
    for seirindex, frac in enumerate(initial_seir_state_distribution):
        state = SEIRState(seirindex + 1)
        k = round(n * frac)
        undefined = [i for i, x in agents_seir_state.items() if x == SEIRState.Undefined]
        if len(undefined) > 0:
            undefined = np.array(undefined)
        for i in np.random.choice(undefined, size=k, replace=False):
            agents_seir_state[i] = state
    for i in [i for i, x in agents_seir_state.items() if x == SEIRState.Undefined]:
        agents_seir_state[i] = SEIRState(random.randint(1, 4))
    return agents_seir_state


def agents_seir_state_get(agents_seir_state, agent_id_or_idx, agents_seir_indices=None):
    return agents_seir_state[agent_id_or_idx]                       # seirstateutil.py:75-76


def agents_seir_state_update(agents_seir_state, new_seir_state, agent_id_or_idx, agents_seir_indices=None):
    agents_seir_state[agent_id_or_idx] = new_seir_state             # seirstateutil.py:89-102
    return agents_seir_state


def update_agent_state(agents_seir_state, agents_infection_type, agents_infection_severity, agentid, agent, agentindex,
                       day, device=0):
    """seirstateutil.py:38-72 for one agent -> 6-tuple or None"""
    sched = agent["state_transition_by_day"]
    if not sched:
        return None
    today = next((e for e in sched if e[0] == day), None)           # which entry fires (reported back to the caller)
    if today is None:
        return None
    cur_state = agents_seir_state[agentid]
    cur_type = agents_infection_type[agentid]                       # KeyError like the reference (:50-54)
    cur_sev = agents_infection_severity[agentid]
    h = _scalar_handle(device)
    packed = np.zeros((1, prm.MAXSCHED), np.int32)
    for j, (d, t, ts) in enumerate(sched):
        packed[0, j] = bridge.pack_sched(d, t, ts)
    h.set_state(seir=np.array([int(cur_state)], np.int8), inf_type=np.array([int(cur_type)], np.int8),
                inf_sev=np.array([int(cur_sev)], np.int8), sched=packed, sched_cnt=np.array([len(sched)], np.uint8))
    h.advance_seir(int(day), None)
    dev = h.get_state()
    new_state, new_type, new_sev = int(dev["seir"][0]), int(dev["inf_type"][0]), int(dev["inf_sev"][0])
    cnt = int(dev["sched_cnt"][0])
    # quirk B1 (seirstateutil.py:118): value 4 in the SEIR slot is what the reference stores for "symptomatic"
    if new_state != int(cur_state):
        agents_seir_state[agentid] = SEIRState(new_state)
    if new_type != int(cur_type):
        agents_infection_type[agentid] = InfectionType(new_type)
    if new_sev != int(cur_sev):
        agents_infection_severity[agentid] = Severity(new_sev)
    agent["state_transition_by_day"] = [bridge.unpack_sched(v) for v in dev["sched"][0, :cnt]]
    return (SEIRState(new_state), SEIRState(int(cur_state)), InfectionType(new_type), Severity(new_sev),
            SEIRStateTransition(int(today[1])), today[2])


def update_agents_state(n_agents, agents_epi, vars_util, agent_ids, day, device=0):
    """batched K4 over `agent_ids` (all agents whose itinerary is generated today) -> ids whose state entry fired"""
    ids = np.asarray(sorted(int(a) for a in agent_ids), np.int64)
    st = bridge.DenseState(n_agents)
    for a in ids:
        e = agents_epi[int(a)]
        ent = e.get("state_transition_by_day") or []
        st.sched_cnt[a] = len(ent)
        for j, (d, t, ts) in enumerate(ent):
            st.sched[a, j] = bridge.pack_sched(d, t, ts)
        st.seir[a] = int(vars_util.agents_seir_state[int(a)])
        st.inf_type[a] = int(vars_util.agents_infection_type.get(int(a), 0))
        st.inf_sev[a] = int(vars_util.agents_infection_severity.get(int(a), 0))
    h = bridge.get_handle(n_agents, 1, 16, prm.build_params(), device)
    before = st.sched_cnt.copy()
    h.set_state(seir=st.seir, inf_type=st.inf_type, inf_sev=st.inf_sev, sched=st.sched, sched_cnt=st.sched_cnt)
    mask = np.zeros(n_agents, np.uint8); mask[ids] = 1
    h.advance_seir(int(day), mask)
    dev = h.get_state()
    fired = [int(a) for a in ids if dev["sched_cnt"][a] != before[a]]
    for a in fired:
        cnt = int(dev["sched_cnt"][a])
        agents_epi[a]["state_transition_by_day"] = [bridge.unpack_sched(v) for v in dev["sched"][a, :cnt]]
        vars_util.agents_seir_state[a] = SEIRState(int(dev["seir"][a]))
        vars_util.agents_infection_type[a] = InfectionType(int(dev["inf_type"][a]))
        vars_util.agents_infection_severity[a] = Severity(int(dev["inf_sev"][a]))
    return fired
