"""The reference-shaped Python API (mtdcovabm_b200.contactnetwork / seirstateutil) driven the way the reference drives
its own classes (contactnetwork_mp.py:287-304, itinerary.py:399): dict containers in, dict containers mutated in place.
Replay mode against the golden days of the unmodified reference."""
from types import SimpleNamespace

import numpy as np
import pytest

from mtdcovabm_b200 import params as prm
from oracle import mtd_oracle as mo
from tests.golden import golden_io as gio

pytestmark = pytest.mark.gpu
NONE = -32768


class FakeStatic:
    """agents_static.get(id, name) (static.py:351-363)"""
    def __init__(self, st):
        self.st = st

    def get(self, a, name):
        return {"age_bracket_index": self.st.age_bracket, "epi_age_bracket_index": self.st.epi_age_bracket,
                "soc_rate": self.st.soc_rate, "res_cellid": self.st.res_cellid}[name][a]


def _rec(row):
    return None if row[0] == NONE else [int(x) for x in row]


def reference_containers(g, st):
    ids = np.nonzero(st.seir != 0)[0]
    epi = {}
    for a in set(ids.tolist()) | set(g.act_ids.tolist()):
        epi[a] = {"state_transition_by_day": [list(mo.unpack_sched(v)) for v in st.sched[a, :st.sched_cnt[a]]],
                  "test_day": _rec(st.test_day[a]), "test_result_day": _rec(st.test_result_day[a]),
                  "quarantine_days": _rec(st.quarantine[a]), "hospitalisation_days": _rec(st.hosp[a]),
                  "vaccination_days": None}
    cat = {}
    for c, a, s, e in zip(g.iv_cell, g.iv_agent, g.iv_start, g.iv_end):
        cat.setdefault(int(c), []).append((int(a), int(s), int(e)))
    vu = SimpleNamespace(cells_agents_timesteps=cat, directcontacts_by_simcelltype_by_day=[],
                         agents_seir_state={int(a): int(st.seir[a]) for a in ids},
                         agents_infection_type={int(a): int(st.inf_type[a]) for a in ids},
                         agents_infection_severity={int(a): int(st.inf_sev[a]) for a in ids},
                         agents_vaccination_doses={int(a): [int(st.last_dose_day[a])] for a in ids
                                                   if st.last_dose_day[a] != NONE})
    return epi, vu


@pytest.mark.parametrize("name,day", list(gio.all_days()))
def test_contactnetwork_dropin_replays_reference_day(name, day):
    from mtdcovabm_b200.contactnetwork import ContactNetwork
    z = gio.load_run(name)
    g = gio.load_day(z, day)
    epi, vu = reference_containers(g, g.pre)
    dyn = SimpleNamespace(masks_hygiene_distancing_multiplier=g.mhd, testing_enabled=g.testing_enabled,
                          quarantine_enabled=g.quarantine_enabled, contact_tracing_enabled=g.contact_tracing_enabled)
    cells_type = {c: int(t) for c, t in enumerate(g.cell_type) if t >= 0}
    indids = {c: int(v) for c, v in enumerate(g.cell_indid) if v >= 0}
    cn = ContactNetwork(g.n_locals, g.N - g.n_locals, 1.0, FakeStatic(g.pre), epi, vu, cells_type, indids, {}, {}, {},
                        prm.load_json("contactnetwork.json"), prm.load_json(gio.RUNS[name]), dyn, mode="replay",
                        tape=g.tape)
    pidx, updated, partial, vu2 = cn.simulate_contact_network(g.day, g.weekday)
    assert pidx == -1 and vu2 is vu
    assert vu.directcontacts_by_simcelltype_by_day == g.contacts.tolist()      # rows, in the reference's order
    assert updated == g.exposed_ids.tolist()
    assert set(partial) == set(updated)
    post = g.post
    for a in updated:
        assert partial[a] is epi[a]
        assert epi[a]["state_transition_by_day"] == [list(mo.unpack_sched(v)) for v in post.sched[a, :post.sched_cnt[a]]]
        assert epi[a]["test_day"] == _rec(post.test_day[a]) and epi[a]["test_result_day"] == _rec(post.test_result_day[a])
        assert epi[a]["hospitalisation_days"] == _rec(post.hosp[a]) and epi[a]["quarantine_days"] == _rec(post.quarantine[a])
        assert (vu.agents_seir_state[a], vu.agents_infection_type[a], vu.agents_infection_severity[a]) == \
            (post.seir[a], post.inf_type[a], post.inf_sev[a])
    # everybody else is untouched
    for a in np.nonzero(post.seir != 0)[0]:
        assert vu.agents_seir_state[int(a)] == post.seir[a]


def test_contactnetwork_per_cell_entry_points_replay_reference_day():
    """simulate_contact_network_by_cellid / generate_contact_network (contactnetwork.py:127-315) cell by cell, in the
    reference's cell order, rebuild the day's rows the way contactnetwork.py:98-105 does -> the reference's rows,
    exposures and final state"""
    from mtdcovabm_b200.contactnetwork import ContactNetwork
    from mtdcovabm_b200 import util
    name, day = "ref_1k_seed6", 16
    z = gio.load_run(name)
    g = gio.load_day(z, day)
    epi, vu = reference_containers(g, g.pre)
    dyn = SimpleNamespace(masks_hygiene_distancing_multiplier=g.mhd, testing_enabled=g.testing_enabled,
                          quarantine_enabled=g.quarantine_enabled, contact_tracing_enabled=g.contact_tracing_enabled)
    cells_type = {c: int(t) for c, t in enumerate(g.cell_type) if t >= 0}
    indids = {c: int(v) for c, v in enumerate(g.cell_indid) if v >= 0}
    cn = ContactNetwork(g.n_locals, g.N - g.n_locals, 1.0, FakeStatic(g.pre), epi, vu, cells_type, indids, {}, {}, {},
                        prm.load_json("contactnetwork.json"), prm.load_json(gio.RUNS[name]), dyn, mode="replay",
                        tape=g.tape)
    rows, exposed = [], []
    first = True
    for cellid in list(vu.cells_agents_timesteps.keys()):
        if first:                                   # generate_contact_network alone: same dict, nothing transmitted
            only, ct = cn.generate_contact_network(cellid)
            first = False
        if len(rows) % 2:       # alternate the two routes: fused per cell / network first, then transmission on the dict
            upd, contacts, ct = cn.simulate_contact_network_by_cellid(cellid, g.day)
        else:
            contacts, ct = cn.generate_contact_network(cellid)
            upd = cn.epi_util.simulate_direct_contacts(contacts, cellid, ct, g.day)
        assert ct == cells_type[cellid]
        sct = int(util.convert_celltype_to_simcelltype(cellid, celltype=ct))
        for (a, b), ov in contacts.items():
            rows.append([sct, a, b, ov[0][0], ov[-1][1]])
        exposed.extend(upd)
    assert rows == g.contacts.tolist()
    assert exposed == g.exposed_ids.tolist() and len(exposed) > 0
    post = g.post
    for a in exposed:
        assert epi[a]["state_transition_by_day"] == [list(mo.unpack_sched(v)) for v in post.sched[a, :post.sched_cnt[a]]]
        assert vu.agents_seir_state[a] == post.seir[a]


def test_update_agent_state_scalar_matches_reference_calls():
    """seirstateutil.update_agent_state, same 7 arguments and return tuple, on calls the reference made"""
    from mtdcovabm_b200 import seirstateutil as ssu
    z = gio.load_run("ref_1k_strict_seed7")
    da = z["adv_day_agent"]; bo, b, ao, a = z["adv_before_off"], z["adv_before"], z["adv_after_off"], z["adv_after"]
    pre, post, applied = z["adv_pre"], z["adv_post"], z["adv_applied"]
    fired = np.nonzero(applied)[0][:40]
    idle = np.nonzero(applied == 0)[0][:10]
    for i in list(fired) + list(idle):
        day, ag = int(da[i, 0]), int(da[i, 1])
        agent = {"state_transition_by_day": [[int(x) for x in e] for e in b[bo[i]:bo[i + 1]]]}
        s, t, v = {ag: int(pre[i, 0])}, {ag: int(pre[i, 1])}, {ag: int(pre[i, 2])}
        out = ssu.update_agent_state(s, t, v, ag, agent, ag, day)
        assert (out is not None) == bool(applied[i])
        assert (s[ag], t[ag], v[ag]) == tuple(int(x) for x in post[i])
        assert agent["state_transition_by_day"] == [[int(x) for x in e] for e in a[ao[i]:ao[i + 1]]]
        if out is not None:
            assert int(out[0]) == int(post[i, 0]) and int(out[1]) == int(pre[i, 0])


@pytest.mark.parametrize("name,ti", list(gio.all_traces()))
def test_epidemiology_contact_tracing_dropin(name, ti):
    """Epidemiology(...).contact_tracing(day) on the reference's containers (sim.py:1929-1968), replay mode"""
    from mtdcovabm_b200.epidemiology import Epidemiology
    z = gio.load_run(name)
    t = gio.load_trace(z, ti)
    N = len(t.epi_present)
    epi = {}
    for a in np.nonzero(t.epi_present)[0]:
        epi[int(a)] = {"state_transition_by_day": [], "quarantine_days": _rec(t.pre.quarantine[a]),
                       "test_day": _rec(t.pre.test_day[a]), "test_result_day": _rec(t.pre.test_result_day[a]),
                       "hospitalisation_days": None}
    # rebuild a contact-row list + start markers whose look-back slices are the recorded ones (the reference drops
    # the last row of every non-current day, quirk B8: one filler row per older day)
    so = t.slice_off
    rows, markers = [], {}
    for back in range(len(so) - 2, -1, -1):
        markers[t.day - back] = len(rows)
        rows.extend([int(r[0]), int(r[1]), int(r[2]), 0, 0] for r in t.rows[so[back]:so[back + 1]])
        if back > 0:
            rows.append([3, int(t.traced[0]), int(t.traced[0]), 0, 0])      # filler: both endpoints traced -> ignored
    vu = SimpleNamespace(contact_tracing_agent_ids=set(int(x) for x in t.traced),
                         directcontacts_by_simcelltype_by_day=rows,
                         directcontacts_by_simcelltype_by_day_start_marker=markers)
    hh = {}
    for c in range(len(t.res_nres)):
        m = t.res_members[t.res_off[c]:t.res_off[c + 1]]
        if len(m):
            hh[c] = {"place": {"resident_uids": np.array(m[:t.res_nres[c]]), "staff_uids": np.array(m[t.res_nres[c]:])}}
    dyn = SimpleNamespace(contact_tracing_enabled=True, testing_enabled=t.testing_enabled,
                          quarantine_enabled=t.quarantine_enabled, statistics=None)
    n_before = len(rows)
    e = Epidemiology(prm.load_json(gio.RUNS[name]), t.n_locals, N - t.n_locals, 1.0, FakeStatic(t.pre), epi, vu, hh, {}, {},
                     dyn, mode="replay", tape=t.tape)
    pidx, updated, agents_epi, vu2 = e.contact_tracing(t.day)
    assert sorted(updated) == t.updated.tolist() and agents_epi is epi
    for a in t.primary:
        assert epi[int(a)]["quarantine_days"] == _rec(t.post.quarantine[a])
        assert epi[int(a)]["test_day"] == _rec(t.post.test_day[a])
        assert epi[int(a)]["test_result_day"] == _rec(t.post.test_result_day[a])
    # clean-up dropped the oldest day and re-based the markers
    assert len(vu.directcontacts_by_simcelltype_by_day) < n_before
    assert min(vu.directcontacts_by_simcelltype_by_day_start_marker.values()) == 0


def test_schedule_vaccinations_dropin_replays_reference_calls():
    """Epidemiology.schedule_vaccinations (epidemiology.py:911-976): six consecutive reference calls
    (tests/golden/make_vacc_kat.py); the dyn_params bookkeeping must match exactly, the reference's own sample is fed
    in replay mode, and in Philox mode the device draws exactly num_vaccinations_today unvaccinated locals"""
    import json, os
    from mtdcovabm_b200.epidemiology import Epidemiology
    K = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "vacc_kat.json")))
    ep = prm.load_json("epidemiology.json")
    for mode in ("replay", "philox"):
        n_locals, n_tourists = K["n_locals"], K["n_tourists"]
        agents_epi = {a: {"vaccination_days": None} for a in range(n_locals + n_tourists)}
        vu = SimpleNamespace(agents_vaccination_doses={})
        dp = SimpleNamespace(vaccination_propensity=0.0, last_vaccination_propensity=0.0, num_agents_to_be_vaccinated=0,
                             num_vaccinations_today=0)
        e = Epidemiology(ep, n_locals, n_tourists, K["ratio"], None, agents_epi, vu, {}, {}, {}, dp, mode=mode, seed=4)
        for row in K["days"]:
            dp.vaccination_propensity = row["propensity"]
            np.random.seed(1000 + row["day"])
            e.tape = (np.array(row["updated"], np.int32), np.array(row["ts"], np.int32))
            upd = e.schedule_vaccinations(row["day"])
            assert dp.num_agents_to_be_vaccinated == row["num_agents_to_be_vaccinated"]
            assert dp.num_vaccinations_today == row["num_vaccinations_today"]
            assert dp.last_vaccination_propensity == row["last_vaccination_propensity"]
            assert len(upd) == len(row["updated"])
            if mode == "replay":
                assert upd == row["updated"]
                assert [agents_epi[a]["vaccination_days"][-1] for a in upd] == [[row["day"] + 1, t] for t in row["ts"]]
            else:
                assert len(set(upd)) == len(upd) and all(a < n_locals and a not in vu.agents_vaccination_doses for a in upd)
                assert all(agents_epi[a]["vaccination_days"][-1][0] == row["day"] + 1 and
                           42 <= agents_epi[a]["vaccination_days"][-1][1] <= 120 for a in upd)
            for a in upd:                                       # the next day's itinerary stage administers the doses
                vu.agents_vaccination_doses.setdefault(a, []).append(row["day"] + 1)

