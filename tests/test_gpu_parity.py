"""GPU parity tests (run on the B200 box): the CUDA path, called through the C ABI, against
 (1) the golden days recorded from the unmodified reference (replay mode: bit-exact rows, exposures, state), and
 (2) the CPU oracle in Philox mode on the golden inputs and on tiled synthetic populations."""
import numpy as np
import pytest

from mtdcovabm_b200 import capi
from mtdcovabm_b200 import params as prm
from oracle import mtd_oracle as mo
from tests import helpers as H
from tests.golden import golden_io as gio

pytestmark = pytest.mark.gpu
DAYS = list(gio.all_days())


@pytest.mark.parametrize("name,day", DAYS)
def test_replay_day_matches_reference(name, day):
    z = gio.load_run(name)
    P = prm.build_params(None, prm.load_json(gio.RUNS[name]))
    g = gio.load_day(z, day)
    h = H.make_handle(P, g.pre, g.cell_type, g.cell_indid, len(g.iv_cell) + 16)
    H.set_tape(h, g.tape)
    d = h.make_day(capi.MODE_REPLAY, g.day, g.mhd, g.testing_enabled, g.quarantine_enabled)
    h.simulate_contact_network(d, g.iv_agent, g.iv_cell, g.iv_start, g.iv_end)
    h.simulate_direct_contacts(d)
    c = h.counters()
    assert c.error == 0 and c.tape_miss == 0 and c.sched_overflow == 0
    rows, dur, cell = h.get_contacts()
    assert np.array_equal(rows, g.contacts)              # the reference's rows, element for element, in order
    assert np.array_equal(h.get_exposed(), g.exposed_ids)
    assert H.state_diff(h.get_state(), g.post) == []
    h.close()


def _oracle_day(P, mode, g_day, mhd, te, qe, seed, s, st):
    ctx = mo.make_ctx(mode, g_day, mhd, te, qe, seed=seed)
    rows, dur, rc, s1 = mo.contact_network(P, ctx, s["iv_cell"], s["iv_agent"], s["iv_start"], s["iv_end"],
                                           s["cell_type"], s["cell_indid"], st)
    exp, p, s2 = mo.transmission(P, ctx, s["iv_cell"], s["iv_agent"], s["iv_start"], s["iv_end"], s["cell_type"], rows,
                                 dur, rc, st, want_p=True)
    return rows, dur, rc, exp, p, s1


@pytest.mark.parametrize("reps,agent_major,seed", [(1, False, 1), (1, True, 2), (8, True, 3), (64, True, 4)])
def test_philox_matches_oracle_on_tiled_population(reps, agent_major, seed):
    z = gio.load_run("ref_1k_seed6")
    P = prm.build_params()
    g = gio.load_day(z, 16)
    s = H.tile_day(g, reps, agent_major=agent_major, seed=seed, infectious_frac=0.05)
    st = s["state"]
    h = H.make_handle(P, st, s["cell_type"], s["cell_indid"], len(s["iv_cell"]))
    h.enable_infection_prob()
    d = h.make_day(capi.MODE_PHILOX, 16, 0.2, True, True, seed=seed)
    h.simulate_contact_network(d, s["iv_agent"], s["iv_cell"], s["iv_start"], s["iv_end"])
    h.simulate_direct_contacts(d)
    c = h.counters()
    assert c.error == 0
    ost = st.copy()
    rows, dur, rc, exp, p, s1 = _oracle_day(P, mo.MODE_PHILOX, 16, 0.2, True, True, seed, s, ost)
    grows, gdur, gcell = h.get_contacts()
    assert c.pairs_evaluated == s1.pairs_evaluated and c.n_potential == s1.n_potential
    assert np.array_equal(grows, rows) and np.array_equal(gdur, dur) and np.array_equal(gcell, rc)
    gp = h.get_infection_prob()
    both = ~np.isnan(p)
    # the sequential oracle skips rows of agents exposed earlier the same day; the GPU evaluates every S-I row
    assert not np.any(np.isnan(gp) & both)
    assert np.allclose(gp[both], p[both], rtol=1e-12, atol=0)       # north_star bar: 1e-6 relative
    assert len(exp) > 0 and np.array_equal(h.get_exposed(), exp)
    assert H.state_diff(h.get_state(), ost) == []
    h.close()


def test_seir_advance_matches_oracle():
    z = gio.load_run("ref_1k_strict_seed7")
    da = z["adv_day_agent"]; bo, b = z["adv_before_off"], z["adv_before"]
    n = len(da)
    st = mo.AgentState(n)
    for i in range(n):
        ent = b[bo[i]:bo[i + 1]]
        st.sched_cnt[i] = len(ent)
        for j, (dd, t, ts) in enumerate(ent):
            st.sched[i, j] = mo.pack_sched(dd, t, ts)
    st.seir[:] = z["adv_pre"][:, 0]; st.inf_type[:] = z["adv_pre"][:, 1]; st.inf_sev[:] = z["adv_pre"][:, 2]
    h = H.make_handle(prm.build_params(), st, np.zeros(4, np.int8), np.full(4, -1, np.int16), 16)
    for day in np.unique(da[:, 0]):
        mask = (da[:, 0] == day).astype(np.uint8)
        h.advance_seir(int(day), mask)
        mo.advance(int(day), st, mask)
    dev = h.get_state()
    assert np.array_equal(np.stack([dev["seir"], dev["inf_type"], dev["inf_sev"]], 1), z["adv_post"])
    assert H.state_diff(dev, st) == []
    assert np.array_equal(h.occupancy(), np.bincount(st.seir, minlength=6)[:6])
    h.close()


def test_two_rank_cell_sharding_equals_single_rank():
    """cells sharded over 2 handles on one device, delta buffers exchanged -> same state as the 1-rank run"""
    import torch
    z = gio.load_run("ref_1k_seed6")
    P = prm.build_params()
    g = gio.load_day(z, 20)
    s = H.tile_day(g, 4, agent_major=True, seed=9, infectious_frac=0.08)
    st = s["state"]
    M = len(s["iv_cell"])
    d_args = (capi.MODE_PHILOX, 20, 0.2, True, True)
    h0 = H.make_handle(P, st, s["cell_type"], s["cell_indid"], M)
    d = h0.make_day(*d_args, seed=5)
    h0.simulate_contact_network(d, s["iv_agent"], s["iv_cell"], s["iv_start"], s["iv_end"])
    h0.simulate_direct_contacts(d)
    ref_state, ref_exp = h0.get_state(), h0.get_exposed()
    ref_rows = h0.get_contacts()[0]
    assert len(ref_exp) > 0
    cap = 4096
    bufs = torch.zeros((2, cap + 1, capi.DELTA_WORDS), dtype=torch.int32, device="cuda")
    hs, nrows = [], 0
    order = np.arange(M, dtype=np.uint32)
    for r in range(2):
        sel = (s["iv_cell"] % 2) == r
        h = H.make_handle(P, st, s["cell_type"], s["cell_indid"], M)
        h.simulate_contact_network(d, np.ascontiguousarray(s["iv_agent"][sel]), np.ascontiguousarray(s["iv_cell"][sel]),
                                   np.ascontiguousarray(s["iv_start"][sel]), np.ascontiguousarray(s["iv_end"][sel]),
                                   iv_order=np.ascontiguousarray(order[sel]))
        h.simulate_direct_contacts(d, delta_buf=bufs[r], cap_records=cap, apply=False)
        h.sync()
        nrows += int(h.counters().n_rows)
        hs.append(h)
    assert nrows == len(ref_rows)
    for h in hs:
        h.apply_deltas(bufs, 2, cap)
        assert H.state_diff(h.get_state(), type("S", (), ref_state | {"sched_cnt": ref_state["sched_cnt"]})) == []
        assert np.array_equal(h.get_exposed(), ref_exp)
        h.close()
    h0.close()


@pytest.mark.parametrize("name,ti", list(gio.all_traces()))
def test_contact_tracing_replay_matches_reference(name, ti):
    """K5 through the C ABI on every contact_tracing call the reference made (epidemiology.py:667-834)"""
    z = gio.load_run(name)
    P = prm.build_params(None, prm.load_json(gio.RUNS[name]))
    t = gio.load_trace(z, ti)
    h = H.make_handle(P, t.pre, np.zeros(len(t.res_nres), np.int8), np.full(len(t.res_nres), -1, np.int16), 16)
    h.set_residences(t.res_off, t.res_nres, t.res_members)
    d = h.make_day(capi.MODE_REPLAY, t.day, 0.0, t.testing_enabled, t.quarantine_enabled)
    h.contact_tracing(d, t.traced, t.slice_off, t.rows, epi_present=t.epi_present, tape=t.tape)
    prim, sec, s = h.get_traced()
    assert s.tape_miss == 0 and s.n_scan_draws == len(t.tape.scan_key)
    assert np.array_equal(np.nonzero(prim)[0], t.primary) and np.array_equal(np.nonzero(sec)[0], t.secondary)
    dev = h.get_state()
    for f in ("quarantine", "test_day", "test_result_day"):
        assert np.array_equal(dev[f], getattr(t.post, f)), f
    h.close()


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_contact_tracing_philox_matches_oracle(seed):
    z = gio.load_run("ref_1k_strict_seed7")
    P = prm.build_params(None, prm.load_json(gio.RUNS["ref_1k_strict_seed7"]))
    for ti in range(len(z["trace_days"])):
        t = gio.load_trace(z, ti)
        traced = np.unique(np.concatenate([t.traced, t.rows[seed::7, 1]])).astype(np.int32)    # more positives
        h = H.make_handle(P, t.pre, np.zeros(len(t.res_nres), np.int8), np.full(len(t.res_nres), -1, np.int16), 16)
        h.set_residences(t.res_off, t.res_nres, t.res_members)
        d = h.make_day(capi.MODE_PHILOX, t.day, 0.0, True, True, seed=seed)
        h.contact_tracing(d, traced, t.slice_off, t.rows, epi_present=t.epi_present)
        prim, sec, s = h.get_traced()
        ost = t.pre.copy()
        ctx = mo.make_ctx(mo.MODE_PHILOX, t.day, 0.0, True, True, seed=seed)
        oprim, osec, os_ = mo.contact_tracing(P, ctx, None, traced, t.slice_off, t.rows, t.res_off, t.res_nres,
                                              t.res_members, ost, t.epi_present, t.n_locals)
        assert s.n_scan_draws == os_.n_scan_draws and prim.sum() > 20
        assert np.array_equal(prim, oprim) and np.array_equal(sec, osec)
        dev = h.get_state()
        for f in ("quarantine", "test_day", "test_result_day"):
            assert np.array_equal(dev[f], getattr(ost, f)), f
        h.close()


def test_pipelined_host_boundary_matches_synchronous_api():
    """engine.step_host (prefetched uploads, results streamed into pinned host memory) == the synchronous calls"""
    import torch
    from mtdcovabm_b200 import engine
    z = gio.load_run("ref_1k_seed6")
    P = prm.build_params()
    days = [gio.load_day(z, d) for d in (12, 16, 20)]
    tiles = [H.tile_day(g, 6, agent_major=True, seed=4, infectious_frac=0.06) for g in days]
    st = tiles[0]["state"]
    N, C = st.n, len(tiles[0]["cell_type"])
    M = max(len(t["iv_cell"]) for t in tiles)

    class Pop:
        pass
    pop = Pop()
    dev = torch.device("cuda", 0)
    for name, arr in (("cell_type", np.asarray(tiles[0]["cell_type"], np.int8)), ("cell_indid", np.asarray(tiles[0]["cell_indid"], np.int16)),
                      ("age_bracket", st.age_bracket), ("epi_age_bracket", st.epi_age_bracket), ("soc_rate", st.soc_rate),
                      ("res_cellid", st.res_cellid), ("seir", st.seir), ("inf_type", st.inf_type), ("inf_sev", st.inf_sev),
                      ("sched", st.sched), ("sched_cnt", st.sched_cnt)):
        setattr(pop, name, torch.from_numpy(np.ascontiguousarray(arr)).to(dev))

    def host(t):
        return {k: torch.from_numpy(t[k]).pin_memory() for k in ("iv_agent", "iv_cell", "iv_start", "iv_end")}
    hs = [host(t) for t in tiles]

    # synchronous reference run through the plain calls
    e0 = engine.DayEngine(N, C, M + 16, P, seed=11)
    e0.set_population(pop)
    want = []
    for k, s in enumerate(hs):
        e0.step(30 + k, s, mhd=0.1, testing=True, quarantine=True, mem=capi.MEM_HOST)
        rows, dur, cell = e0.h.get_contacts()
        want.append((rows, e0.h.get_exposed(), e0.h.occupancy(), int(e0.h.counters().pairs_evaluated)))
    final0 = e0.h.get_state()
    e0.close()

    e1 = engine.DayEngine(N, C, M + 16, P, seed=11)
    e1.set_population(pop)
    e1.enable_host_pipeline(max(len(w[0]) for w in want) + 64)
    got = []
    for k, s in enumerate(hs):
        nxt = hs[k + 1] if k + 1 < len(hs) else None
        r = e1.step_host(30 + k, s, next_s=nxt, mhd=0.1, testing=True, quarantine=True)
        # copy out at once: the pinned buffers are reused two days later
        got.append((r.rows.copy(), r.exposed().copy(), r.occupancy.copy(), r.pairs_evaluated))
    final1 = e1.h.get_state()
    e1.close()
    assert sum(len(w[1]) for w in want) > 0
    for (rows, exp, occ, pairs), (r12, e12, o12, p12) in zip(want, got):
        assert len(r12) == len(rows)
        assert np.array_equal(r12["simcelltype"], rows[:, 0]) and np.array_equal(r12["a1"], rows[:, 1])
        assert np.array_equal(r12["a2"], rows[:, 2]) and np.array_equal(r12["start"], rows[:, 3])
        assert np.array_equal(r12["end"], rows[:, 4])
        assert np.array_equal(e12, exp)
        assert np.array_equal(o12, occ) and p12 == pairs
    for f in H.STATE_FIELDS:
        assert np.array_equal(final0[f], final1[f]), f


def test_intervention_mask_matches_oracle():
    """K0's intervention mask (deceased / quarantined / hospitalised agents keep only their residence-cell intervals;
    SURVEY.md section 8f rank 1) == the oracle run on the stream filtered by mtdo_mask_intervals"""
    P = prm.build_params()
    g = gio.load_day(gio.load_run("ref_1k_seed6"), 16)
    s = H.tile_day(g, 6, agent_major=True, seed=4, infectious_frac=0.05)
    st = s["state"]
    rng = np.random.default_rng(8)
    ids = np.unique(s["iv_agent"])
    day = 20
    pick = rng.choice(ids, size=len(ids) // 4, replace=False)
    q, hsp, dec, edge = np.array_split(pick, 4)
    st.quarantine[q] = np.stack([np.full(len(q), day - 2), np.full(len(q), 50), np.full(len(q), day + 3)], 1)
    st.hosp[hsp] = np.stack([np.full(len(hsp), day), np.full(len(hsp), 60), np.full(len(hsp), day)], 1)
    st.seir[dec] = 5
    st.quarantine[edge] = np.stack([np.full(len(edge), day + 1), np.full(len(edge), 50), np.full(len(edge), day + 9)], 1)  # not yet
    keep = mo.mask_intervals(day, s["iv_agent"], s["iv_cell"], st)
    assert 0 < keep.sum() < len(keep)
    assert keep[np.isin(s["iv_agent"], edge)].all()

    h = H.make_handle(P, st, s["cell_type"], s["cell_indid"], len(s["iv_cell"]))
    h.set_intervention_mask(True)
    d = h.make_day(capi.MODE_PHILOX, day, 0.1, True, True, seed=9)
    h.simulate_contact_network(d, s["iv_agent"], s["iv_cell"], s["iv_start"], s["iv_end"])
    h.simulate_direct_contacts(d)
    c = h.counters()
    assert c.error == 0
    grow, gdur, gcell = h.get_contacts()

    ost = st.copy()
    ctx = mo.make_ctx(mo.MODE_PHILOX, day, 0.1, True, True, seed=9)
    f = [np.ascontiguousarray(s[k][keep]) for k in ("iv_cell", "iv_agent", "iv_start", "iv_end")]
    orow, odur, orc, s1 = mo.contact_network(P, ctx, f[0], f[1], f[2], f[3], s["cell_type"], s["cell_indid"], ost)
    oexp, _, _ = mo.transmission(P, ctx, f[0], f[1], f[2], f[3], s["cell_type"], orow, odur, orc, ost)
    assert np.array_equal(grow, orow) and np.array_equal(gdur, odur) and np.array_equal(gcell, orc)
    assert c.pairs_evaluated == s1.pairs_evaluated
    assert np.array_equal(h.get_exposed(), oexp)
    assert H.state_diff(h.get_state(), ost) == []
    # masked agents meet nobody outside their residence cell
    masked = np.concatenate([q, hsp, dec])
    res = st.res_cellid
    for col in (1, 2):
        m = np.isin(grow[:, col], masked)
        assert np.all(gcell[m] == res[grow[m, col]])
    h.set_intervention_mask(False)
    h.close()


@pytest.mark.parametrize("quarantine_enabled", [True, False])
def test_test_result_processing_matches_oracle(quarantine_enabled):
    """K6 (itinerary.py:2301-2332): positives, Positive quarantines and cleared result days == oracle, bit for bit"""
    P = prm.build_params(None, prm.load_json("epidemiology_strict_days.json"))
    n, day = 40_000, 23
    rng = np.random.default_rng(12)
    st = mo.AgentState(n)
    st.seir[:] = rng.integers(0, 6, n)
    due = rng.random(n) < 0.3
    st.test_result_day[due, 0] = day
    st.test_result_day[due, 1] = rng.integers(42, 121, due.sum())
    past = ~due & (rng.random(n) < 0.2)
    st.test_result_day[past, 0] = day - rng.integers(1, 5, past.sum()); st.test_result_day[past, 1] = 77
    future = ~due & ~past & (rng.random(n) < 0.2)
    st.test_result_day[future, 0] = day + 2; st.test_result_day[future, 1] = 50
    hasq = rng.random(n) < 0.3                         # existing quarantines: earlier start blocks, later start is replaced
    st.quarantine[hasq, 0] = day + rng.integers(-3, 4, hasq.sum()); st.quarantine[hasq, 1] = 60
    st.quarantine[hasq, 2] = st.quarantine[hasq, 0] + 10
    present = (rng.random(n) < 0.95).astype(np.uint8)
    h = capi.Handle(n, 1, 16)
    h.set_params(P)
    H.upload_state(h, st)
    d = h.make_day(capi.MODE_PHILOX, day, 0.0, True, quarantine_enabled, seed=21)
    got = h.process_test_results(d, present)
    ost = st.copy()
    ctx = mo.make_ctx(mo.MODE_PHILOX, day, 0.0, True, quarantine_enabled, seed=21)
    exp = mo.process_test_results(P, ctx, ost, present)
    assert len(exp) > 100 and np.array_equal(got, exp)
    assert H.state_diff(h.get_state(), ost) == []
    # rates: ~ (1 - false negative) of the due Exposed/Infectious, ~ false positive of the others
    ei = due & (present == 1) & np.isin(st.seir, (2, 3))
    oth = due & (present == 1) & ~np.isin(st.seir, (2, 3))
    flag = np.zeros(n, bool); flag[got] = True
    assert abs(flag[ei].mean() - (1 - P.testing_false_negative_rate)) < 0.03
    assert abs(flag[oth].mean() - P.testing_false_positive_rate) < 0.02
    assert not flag[~due].any()
    if not quarantine_enabled:
        assert np.array_equal(ost.quarantine, st.quarantine)
    h.close()


@pytest.mark.parametrize("k", [0, 1, 777, 30_000, 10**9])
def test_vaccination_scheduling_matches_oracle(k):
    """K7 (epidemiology.py:911-976 device part + itinerary.py:2489-2527): exact-k radix select == oracle's sorted keys;
    dose bookkeeping on the vaccination day through mtd_advance_seir"""
    n, n_locals, day = 60_000, 50_000, 31
    rng = np.random.default_rng(5)
    last = np.full(n, prm.NONE, np.int32)
    done = rng.random(n) < 0.4
    last[done] = rng.integers(1, day, done.sum())
    h = capi.Handle(n, 1, 16)
    h.set_params(prm.build_params())
    h.enable_vaccination(n_locals)
    h.set_vaccination(last, np.full((n, 2), prm.NONE, np.int32))
    d = h.make_day(capi.MODE_PHILOX, day, 0.0, False, False, seed=17)
    got = h.schedule_vaccinations(d, k)
    o_last, o_rec = last.copy(), np.full((n, 2), prm.NONE, np.int32)
    ctx = mo.make_ctx(mo.MODE_PHILOX, day, 0.0, False, False, seed=17)
    exp = mo.schedule_vaccinations(ctx, n_locals, k, o_last, o_rec)
    n_cand = int((last[:n_locals] == prm.NONE).sum())
    assert len(exp) == min(k, n_cand) and np.array_equal(got, exp)
    g_last, g_rec = h.get_vaccination()
    assert np.array_equal(g_rec, o_rec) and np.array_equal(g_last, o_last)
    if len(exp):
        assert np.all(g_rec[exp, 0] == day + 1) and g_rec[exp, 1].min() >= 42 and g_rec[exp, 1].max() <= 120
        assert np.all(last[exp] == prm.NONE) and exp.max() < n_locals
    # next day: the dose lands; the day after: the record is cleared
    for dd in (day + 1, day + 2):
        h.advance_seir(dd, None)
        mo.vaccination_day(dd, n_locals, o_last, o_rec)
        g_last, g_rec = h.get_vaccination()
        assert np.array_equal(g_rec, o_rec) and np.array_equal(g_last, o_last)
    if len(exp):
        assert np.all(g_last[exp] == day + 1) and np.all(g_rec[exp, 0] == prm.NONE)
    h.close()


def test_vaccination_sample_is_uniform():
    """the k-subset is uniform over the candidates: per-agent selection frequency over many days ~ k / n_candidates"""
    n, k = 2000, 100
    h = capi.Handle(n, 1, 16)
    h.set_params(prm.build_params())
    h.enable_vaccination(n)
    cnt = np.zeros(n)
    for day in range(1, 401):
        h.set_vaccination(np.full(n, prm.NONE, np.int32), np.full((n, 2), prm.NONE, np.int32))
        ids = h.schedule_vaccinations(h.make_day(capi.MODE_PHILOX, day, 0.0, False, False, seed=3), k)
        assert len(ids) == k
        cnt[ids] += 1
    p = k / n
    z = (cnt - 400 * p) / np.sqrt(400 * p * (1 - p))
    assert abs(z.mean()) < 0.2 and 0.8 < z.std() < 1.2 and np.abs(z).max() < 5
    h.close()


def test_checkpoint_resume_is_bit_exact(tmp_path):
    """8 days straight == 4 days, checkpoint, new engine, resume, 4 days (state, exposures)"""
    import torch
    from mtdcovabm_b200 import engine, synth
    dev = torch.device("cuda", 0)
    bank = synth.load_bank()
    pop = synth.SynthPopulation(bank, 12, dev, seed=2, infectious_frac=0.05)
    P = prm.build_params()
    streams = [pop.stream(k) for k in range(pop.n_days)]
    max_m = max(len(s["iv_agent"]) for s in streams) + 16
    n_locals = pop.n_agents

    def fresh():
        e = engine.DayEngine(pop.n_agents, pop.n_cells, max_m, P, device=0, seed=9)
        e.set_population(pop)
        e.enable_vaccination(n_locals)
        return e

    def run(e, first, last):
        out = []
        for d in range(first, last + 1):
            e.step(d, streams[d % pop.n_days], testing=True, quarantine=True)
            out.append(int(e.h.counters().n_exposed))
            if d % 3 == 0:
                e.h.schedule_vaccinations(e.h.make_day(capi.MODE_PHILOX, d, 0.0, seed=9), 500)
        return out

    a = fresh()
    exp_a = run(a, 1, 8)
    st_a, vac_a = a.h.get_state(), a.h.get_vaccination()
    a.close()
    b = fresh()
    exp_b = run(b, 1, 4)
    ck = str(tmp_path / "day4.npz")
    b.save_checkpoint(ck, 4)
    b.close()
    c = fresh()
    assert c.load_checkpoint(ck) == 4
    exp_b += run(c, 5, 8)
    st_c, vac_c = c.h.get_state(), c.h.get_vaccination()
    c.close()
    assert exp_a == exp_b and sum(exp_a) > 0
    for k in st_a:
        assert np.array_equal(st_a[k], st_c[k]), k
    assert np.array_equal(vac_a[0], vac_c[0]) and np.array_equal(vac_a[1], vac_c[1]) and (vac_a[0] != prm.NONE).sum() > 0


def test_k0_overlap_equals_serial_day_loop():
    """mtd_set_k0_overlap: K0 of day d+1 on its own stream beside K1..K3 of day d (double-buffered K0 products) gives
    the same rows, exposures and final state as the serial order, over more days than there are buffers"""
    import torch
    from mtdcovabm_b200 import engine
    z = gio.load_run("ref_1k_seed6")
    P = prm.build_params()
    days = [gio.load_day(z, d) for d in (12, 16, 20)]
    tiles = [H.tile_day(g, 6, agent_major=True, seed=4, infectious_frac=0.06) for g in days]
    st = tiles[0]["state"]
    N, C = st.n, len(tiles[0]["cell_type"])
    M = max(len(t["iv_cell"]) for t in tiles)

    class Pop:
        pass
    pop = Pop()
    dev = torch.device("cuda", 0)
    for name, arr in (("cell_type", np.asarray(tiles[0]["cell_type"], np.int8)), ("cell_indid", np.asarray(tiles[0]["cell_indid"], np.int16)),
                      ("age_bracket", st.age_bracket), ("epi_age_bracket", st.epi_age_bracket), ("soc_rate", st.soc_rate),
                      ("res_cellid", st.res_cellid), ("seir", st.seir), ("inf_type", st.inf_type), ("inf_sev", st.inf_sev),
                      ("sched", st.sched), ("sched_cnt", st.sched_cnt)):
        setattr(pop, name, torch.from_numpy(np.ascontiguousarray(arr)).to(dev))
    ds = [{k: torch.from_numpy(np.ascontiguousarray(t[k])).to(dev) for k in ("iv_agent", "iv_cell", "iv_start", "iv_end")}
          for t in tiles]
    # an empty day in the middle: it takes the serial path (nothing to overlap) between two overlapped days
    ds.insert(2, {"iv_agent": torch.empty(0, dtype=torch.int32, device=dev), "iv_cell": torch.empty(0, dtype=torch.int32, device=dev),
                  "iv_start": torch.empty(0, dtype=torch.uint8, device=dev), "iv_end": torch.empty(0, dtype=torch.uint8, device=dev)})
    torch.cuda.synchronize()
    n_days = 9

    def run(overlap, read_every_day):
        e = engine.DayEngine(N, C, M + 16, P, seed=11)
        e.set_population(pop)
        e.set_inputs_ready(overlap)
        out = []
        for k in range(n_days):
            e.step(30 + k, ds[k % len(ds)], mhd=0.1, testing=True, quarantine=True)
            if read_every_day:
                rows, dur, cell = e.h.get_contacts()
                out.append((rows, dur, e.h.get_exposed(), int(e.h.counters().pairs_evaluated)))
        rows, dur, cell = e.h.get_contacts()
        last = (rows, dur, e.h.get_exposed(), e.h.occupancy(), e.h.counters())
        final = e.h.get_state()
        e.close()
        return out, last, final

    want, wlast, wfinal = run(False, True)
    got, glast, gfinal = run(True, True)
    assert sum(len(w[2]) for w in want) > 0
    for (r0, d0, e0, p0), (r1, d1, e1, p1) in zip(want, got):
        assert np.array_equal(r0, r1) and np.array_equal(d0, d1) and np.array_equal(e0, e1) and p0 == p1
    # without host reads between the days the host runs ahead and the stages really overlap
    _, alast, afinal = run(True, False)
    assert np.array_equal(wlast[0], alast[0]) and np.array_equal(wlast[1], alast[1]) and np.array_equal(wlast[2], alast[2])
    assert np.array_equal(wlast[3], alast[3])
    assert int(alast[4].error) == 0 and int(alast[4].n_rows) == int(wlast[4].n_rows)
    for f in H.STATE_FIELDS:
        assert np.array_equal(wfinal[f], gfinal[f]), f
        assert np.array_equal(wfinal[f], afinal[f]), f
