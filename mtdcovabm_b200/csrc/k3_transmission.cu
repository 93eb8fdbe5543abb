// K3 -- transmission (Epidemiology.simulate_direct_contacts, epidemiology.py:88-231), whole-trajectory sampling
// at exposure time (simulate_seir_state_transition, epidemiology.py:233-500) and the scheduling it triggers
// (schedule_hospitalisation :647-660, schedule_test :502-565, schedule_quarantine :568-625).
// K4 -- daily SEIR advance (seirstateutil.update_agent_state, seirstateutil.py:38-72, table :105-136).
//
// The reference walks the day's contact rows sequentially; a newly exposed agent is not susceptible for later rows
// (epidemiology.py:94,221).  Here every S-I row is evaluated independently with a draw keyed by the row, and the
// earliest successful row (atomicMin on an order key) wins -- identical to the sequential result because draws are
// keyed, not positional (SURVEY.md section 7 "within-day visibility").  The winners' trajectories are written as
// fixed-size delta records straight into the buffer that is exchanged between GPUs; every rank then applies all
// records to its replicated per-agent state.
#include "common.cuh"

namespace {

__device__ __forceinline__ void vacc_multipliers(const mtd_params *P, int last, int day, double &imm, double &asym)
{   // util.calculate_vaccination_multipliers, util.py:117-150
    imm = 1.0; asym = 1.0;
    if (last == MTD_NONE) return;
    int d = day - last;
    if (d == 0) { imm = 1.0 - P->vacc_immunity; asym = 1.0 - P->vacc_asymptomatic; return; }
    double x = (double)d / P->vacc_decay_interval;
    if (x < 1) x = 1;
    imm = 1.0 - pow(P->vacc_immunity, x);
    asym = 1.0 - pow(P->vacc_asymptomatic, x);
}

__device__ __forceinline__ long long tape_find_pair(const DevTape &T, int cell, int a1, int a2)
{
    uint64_t k1 = ((uint64_t)(uint32_t)cell << 32) | (uint32_t)a1;
    uint32_t k2 = (uint32_t)a2;
    long long lo = 0, hi = T.n_pair;
    while (lo < hi) {
        long long mid = (lo + hi) >> 1;
        if (T.pair_k1[mid] < k1 || (T.pair_k1[mid] == k1 && T.pair_k2[mid] < k2)) lo = mid + 1; else hi = mid;
    }
    return (lo < T.n_pair && T.pair_k1[lo] == k1 && T.pair_k2[lo] == k2) ? lo : -1;
}

__device__ __forceinline__ long long tape_find_expo(const DevTape &T, int agent)
{
    long long lo = 0, hi = T.n_expo;
    while (lo < hi) { long long mid = (lo + hi) >> 1; if (T.expo_agent[mid] < agent) lo = mid + 1; else hi = mid; }
    return (lo < T.n_expo && T.expo_agent[lo] == agent) ? lo : -1;
}

__device__ __forceinline__ unsigned long long row_key(const DevCtx &c, const RowRec &r, int64_t i)
{
    return ((unsigned long long)c.slot_first[r.slot] << 32) | (unsigned long long)(uint32_t)i;
}

// one thread per dense contact row
__global__ void k3_eval_kernel(DevCtx c)
{
    const int64_t n = c.ctr->n_rows;
    const mtd_params *P = c.P;
    int n_si = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        RowRec r = c.rows[i];
        int s1 = c.seir[r.a1], s2 = c.seir[r.a2];
        bool si = (s1 == S_INF && s2 == S_SUS) || (s1 == S_SUS && s2 == S_INF);        // epidemiology.py:106-107
        int dur = (int)(r.sed >> 16);
        if (!si || dur <= 0) { if (c.row_p) c.row_p[i] = __longlong_as_double(0x7ff8000000000000LL); continue; }
        n_si++;
        int ex = (s1 == S_SUS) ? r.a1 : r.a2;
        int cell = c.slot_cell[r.slot];
        int sct = simcelltype_of(c.cell_type[cell]);
        double base;                                                                    // epidemiology.py:998-1014
        if (sct == SCT_RESIDENCE) base = P->p_household;
        else if (sct == SCT_WORKPLACE) base = (cell == c.res_cellid[ex]) ? P->p_workplace : P->p_community;
        else if (sct == SCT_SCHOOL) base = P->p_school;
        else base = P->p_community;
        double lg = (dur >= 0 && dur < MTD_LOGTAB_N) ? c.logtab[dur] : log((double)dur);   // math.log of an integer: glibc table
        double mult = lg > 1.0 ? lg : 1.0;                                              // max(1, math.log(duration))
        double imm, asym;
        vacc_multipliers(P, c.last_dose_day[ex], c.day.day, imm, asym);
        double susc = P->probs_by_age[0][c.epi_age_bracket[ex]];
        double p = base * (mult * susc * (1 - c.day.mhd) * imm);                        // epidemiology.py:196
        if (c.row_p) c.row_p[i] = p;
        double u;
        if (c.day.mode == MTD_MODE_REPLAY) {
            long long t = tape_find_pair(c.tape, cell, r.a1, r.a2);
            if (t < 0) continue;                 // the reference never evaluated this row (agent already exposed)
            u = c.tape.pair_u[t];
        } else {
            uint32_t w[4];
            philox4x32((uint32_t)c.day.day, (uint32_t)cell, (uint32_t)r.a1, (uint32_t)r.a2, c.day.seed, SITE_PAIR, w);
            u = u53(w[0], w[1]);
        }
        if (u < p) {
            atomicMin(&c.win[ex], row_key(c, r, i));
            int idx = atomicAdd(&c.ctr->n_succ, 1);
            c.succ[idx] = (int32_t)i;
        }
    }
    n_si = __reduce_add_sync(FULLMASK, n_si);            // whole warps reach this point (no early return above)
    if ((threadIdx.x & 31) == 0 && n_si) atomicAdd(&c.ctr->n_si_rows, n_si);
}

// Draw slots of one exposure (the layout of a replay-tape exposure record, tests/golden/ref_harness.py): 0 exposure
// timestep, 1..10 lognormal day counts of transitions 0..9, 11..14 branch uniforms, 15 / 19 timestep options, 16 / 20
// uniforms, 17 / 18 test / result delays.  Replay mode reads the reference's values; Philox mode fills the same record
// with counter-based draws keyed by (day, agent, slot) -- one slot per lane, all at once -- and reads it the same way.
struct Draws {
    const double *row; int miss;
    __device__ double uniform(int slot)
    {
        if (!row || isnan(row[slot])) { miss++; return 1.0; }
        return row[slot];
    }
    __device__ int lognormal_days(int slot)
    {   // util.sample_log_normal(..., isInt=True): np.random.lognormal then round (util.py:88-115)
        if (!row || isnan(row[slot])) { miss++; return 0; }
        return (int)row[slot];
    }
    __device__ int timestep_option(int slot)
    {   // np.random.choice(np.arange(42, 121)) (epidemiology.py:84)
        if (!row || isnan(row[slot])) { miss++; return 42; }
        return (int)row[slot];
    }
};

__device__ __forceinline__ double philox_slot_draw(const DevCtx &c, const mtd_params *P, int agent, int slot)
{
    uint32_t w[4];
    philox4x32((uint32_t)c.day.day, (uint32_t)agent, (uint32_t)slot, 0u, c.day.seed, SITE_EXPO, w);
    if ((slot >= 1 && slot <= 10) || slot == 17 || slot == 18) {
        double mu, sigma;
        if (slot <= 10) { mu = P->ln_mu[slot - 1]; sigma = P->ln_sigma[slot - 1]; }
        else if (slot == 17) { mu = P->test_mu; sigma = P->test_sigma; }
        else { mu = P->result_mu; sigma = P->result_sigma; }
        double u1 = u53(w[0], w[1]), u2 = u53(w[2], w[3]);
        double z = sqrt(-2.0 * log(1.0 - u1)) * cos(6.283185307179586 * u2);
        return rint(exp(mu + sigma * z));
    }
    if (slot == 15 || slot == 19) return (double)(42 + (int)__umulhi(w[0], 79u));
    return u53(w[0], w[1]);
}


// simulate_seir_state_transition + schedule_test / schedule_quarantine as reached from it (epidemiology.py:233-500):
// the whole future trajectory of one exposure, sampled at once.  `incr0` is the exposure day (today for a contact of the
// day step; a day before the start of the simulation for the itinerary's initial cases, itinerary.py:373-395); entries
// dated before `current_day` are dropped (:255,280,...) and `first_T` remembers the first one kept, which decides the
// states of an agent exposed before day 1 (reset_states, :243).
struct Traj {
    int ns, inf_type, first_T, recovered, symptomatic_day, start_hosp, end_hosp, test_day, test_ts, tres_day;
    int32_t sched[6];
};

__device__ __forceinline__ void sample_trajectory(const mtd_params *P, Draws &D, int b, int current_day, int incr0, int ts,
                                                  double asym, bool testing_enabled, bool quarantine_enabled, bool has_test,
                                                  Traj &t)
{
    int incr = incr0, ns = 0, first_T = -1;
#define STEP(T) do { incr += D.lognormal_days(1 + (T)); \
                     if (incr >= current_day && ns < 6) { if (ns == 0) first_T = (T); \
                         t.sched[ns++] = (int32_t)(((uint32_t)incr << 16) | ((uint32_t)(T) << 8) | (uint32_t)ts); } } while (0)
    int symptomatic_day = -1, recovered = 0, start_hosp = MTD_NONE, end_hosp = MTD_NONE, inf_type;
    STEP(T_EXP_INF);
    if (D.uniform(11) < P->probs_by_age[1][b] * asym) {
        inf_type = IT_PRESYMP;
        STEP(T_INF_SYMP); symptomatic_day = incr;
        if (D.uniform(12) < P->probs_by_age[2][b]) {
            STEP(T_SYMP_SEV); start_hosp = incr;
            if (D.uniform(13) < P->probs_by_age[3][b]) {
                STEP(T_SEV_CRI);
                if (D.uniform(14) < P->probs_by_age[4][b]) { STEP(T_CRI_DEATH); end_hosp = incr; }
                else { STEP(T_CRI_REC); end_hosp = incr; recovered = 1; }
            } else { STEP(T_SEV_REC); end_hosp = incr; recovered = 1; }
        } else { STEP(T_MILD_REC); recovered = 1; }
    } else {
        inf_type = IT_PREASYMP;
        STEP(T_ASYMP_REC); recovered = 1;
    }
    if (recovered) STEP(T_REC_SUS);
#undef STEP
    int test_day = MTD_NONE, test_ts = 0, tres_day = MTD_NONE;
    if (symptomatic_day > -1) {
        int tts = D.timestep_option(15);                                            // epidemiology.py:480
        if (testing_enabled && !has_test) {                                         // schedule_test :502-565
            if (D.uniform(16) < P->testing_after_symptoms_probability) {
                test_day = symptomatic_day + D.lognormal_days(17);
                test_ts = tts;
                tres_day = test_day + D.lognormal_days(18);
            }
        }
        (void)D.timestep_option(19);                                                // epidemiology.py:496
        // schedule_quarantine(Symptomatic) draws one uniform; its result is overwritten by the caller with the
        // value read before the call (epidemiology.py:206,216; itinerary.py:390) -> no state change, only the draw
        // is consumed.
        if (quarantine_enabled) (void)D.uniform(20);
    }
    t.ns = ns; t.inf_type = inf_type; t.first_T = first_T; t.recovered = recovered; t.symptomatic_day = symptomatic_day;
    t.start_hosp = start_hosp; t.end_hosp = end_hosp; t.test_day = test_day; t.test_ts = test_ts; t.tres_day = tres_day;
}

#define K3T_WARPS 4
#define K3T_MAXIV 16
// one WARP per successful draw; only the winner of each agent samples the trajectory
__global__ void __launch_bounds__(K3T_WARPS * 32) k3_trajectory_kernel(DevCtx c, int32_t *dbuf, int64_t cap)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) *c.n_succ_pub = c.ctr->n_succ;     // final: the eval kernel has finished
    __shared__ double s_row[K3T_WARPS][MTD_EXPO_W + 3];
    __shared__ uint16_t s_iv[K3T_WARPS][2][K3T_MAXIV];
    __shared__ uint32_t s_ord[K3T_WARPS][2][K3T_MAXIV];
    const int n = c.ctr->n_succ;
    const mtd_params *P = c.P;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_warps = gridDim.x * K3T_WARPS;
    for (int t = blockIdx.x * K3T_WARPS + warp; t < n; t += n_warps) {
        const int64_t i = c.succ[t];
        RowRec r = c.rows[i];
        const int ex = (c.seir[r.a1] == S_SUS) ? r.a1 : r.a2;
        const unsigned long long key = row_key(c, r, i);
        if (c.win[ex] != key) continue;                      // warp-uniform
        Draws D; D.row = nullptr; D.miss = 0;
        int ts = 0;
        __syncwarp();
        if (c.day.mode == MTD_MODE_REPLAY) {
            long long e = tape_find_expo(c.tape, ex);
            if (e >= 0) D.row = c.tape.expo_row + e * MTD_EXPO_W; else D.miss++;
            // sampled_exposed_timestep = np.random.choice(overlapping_timesteps) (epidemiology.py:246)
            ts = (D.row && !isnan(D.row[0])) ? (int)D.row[0] : 0;
        } else if (c.ov_off) {
            // contacts handed in by the caller (Epidemiology.simulate_direct_contacts on a dict of overlap lists):
            // the flattened timestep list is the row's own overlap list (epidemiology.py:171-174)
            if (lane >= 1 && lane < MTD_EXPO_W) s_row[warp][lane] = philox_slot_draw(c, P, ex, lane);
            __syncwarp();
            if (lane == 0) {
                uint32_t w[4];
                philox4x32((uint32_t)c.day.day, (uint32_t)ex, 0u, 0u, c.day.seed, SITE_EXPO, w);
                int idx = (int)__umulhi(w[0], (uint32_t)(r.sed >> 16));
                for (int64_t k = c.ov_off[i]; k < c.ov_off[i + 1]; k++) {
                    const int lo = c.ov_se[2 * k], len = (int)c.ov_se[2 * k + 1] - lo + 1;
                    if (idx < len) { ts = lo + idx; break; }
                    idx -= len;
                }
            }
            D.row = s_row[warp];
        } else {
            if (lane >= 1 && lane < MTD_EXPO_W) s_row[warp][lane] = philox_slot_draw(c, P, ex, lane);
            // the two agents' intervals in list (emission) order: lanes scan the cell's records as K0 laid them out
            // (arrival order of the scatter), gather the two agents' records and order those few by emission index
            const int64_t b = c.slot_off[r.slot];
            const int m = c.slot_m[r.slot];
            int na = 0, nb = 0;
            for (int p0 = 0; p0 < m; p0 += 32) {
                const int p = p0 + lane;
                IvRec rec; rec.agent = -1; rec.se = 0; rec.order = 0;
                if (p < m) rec = c.csr[b + p];
                const bool ia = p < m && rec.agent == r.a1, ib = p < m && rec.agent == r.a2;
                const uint32_t ba = __ballot_sync(FULLMASK, ia), bb = __ballot_sync(FULLMASK, ib);
                if (ia) { int k = na + __popc(ba & lanemask_lt());
                          if (k < K3T_MAXIV) { s_iv[warp][0][k] = (uint16_t)(rec.se & 0xffffu); s_ord[warp][0][k] = rec.order; } }
                if (ib) { int k = nb + __popc(bb & lanemask_lt());
                          if (k < K3T_MAXIV) { s_iv[warp][1][k] = (uint16_t)(rec.se & 0xffffu); s_ord[warp][1][k] = rec.order; } }
                na += __popc(ba); nb += __popc(bb);
            }
            __syncwarp();
            if (lane == 0) {
                uint32_t w[4];
                philox4x32((uint32_t)c.day.day, (uint32_t)ex, 0u, 0u, c.day.seed, SITE_EXPO, w);
                int idx = (int)__umulhi(w[0], (uint32_t)(r.sed >> 16));
                // flattened overlap list of key (a1, a2): a1's intervals (outer) x a2's (inner), list order
                bool done = false;
                if (na <= K3T_MAXIV && nb <= K3T_MAXIV) {
                    for (int side = 0; side < 2; side++) {       // insertion sort by emission index (a handful of entries)
                        const int cnt = side ? nb : na;
                        for (int x = 1; x < cnt; x++) {
                            const uint32_t o = s_ord[warp][side][x]; const uint16_t v = s_iv[warp][side][x];
                            int y = x - 1;
                            while (y >= 0 && s_ord[warp][side][y] > o) {
                                s_ord[warp][side][y + 1] = s_ord[warp][side][y]; s_iv[warp][side][y + 1] = s_iv[warp][side][y]; y--;
                            }
                            s_ord[warp][side][y + 1] = o; s_iv[warp][side][y + 1] = v;
                        }
                    }
                    for (int p = 0; p < na && !done; p++) {
                        const int sa = s_iv[warp][0][p] & 0xff, ea = s_iv[warp][0][p] >> 8;
                        for (int q = 0; q < nb; q++) {
                            const int lo = max(sa, (int)(s_iv[warp][1][q] & 0xff)), hi = min(ea, (int)(s_iv[warp][1][q] >> 8));
                            if (lo > hi) continue;
                            const int len = hi - lo + 1;
                            if (idx < len) { ts = lo + idx; done = true; break; }
                            idx -= len;
                        }
                    }
                } else {      // more intervals than the staging holds: walk the records in emission order by repeated
                              // minimum search (emission indices are unique)
                    long long la = -1;
                    for (int p = 0; p < na && !done; p++) {
                        uint32_t best = 0xffffffffu; int sa = 0, ea = -1;
                        for (int x = 0; x < m; x++) {
                            const IvRec ra = c.csr[b + x];
                            if (ra.agent == r.a1 && (long long)ra.order > la && ra.order <= best) { best = ra.order; sa = ra.se & 0xff; ea = (ra.se >> 8) & 0xff; }
                        }
                        la = (long long)best;
                        long long lb = -1;
                        for (int q = 0; q < nb; q++) {
                            uint32_t bestb = 0xffffffffu; int sb = 0, eb = -1;
                            for (int x = 0; x < m; x++) {
                                const IvRec rb = c.csr[b + x];
                                if (rb.agent == r.a2 && (long long)rb.order > lb && rb.order <= bestb) { bestb = rb.order; sb = rb.se & 0xff; eb = (rb.se >> 8) & 0xff; }
                            }
                            lb = (long long)bestb;
                            const int lo = max(sa, sb), hi = min(ea, eb);
                            if (lo > hi) continue;
                            const int len = hi - lo + 1;
                            if (idx < len) { ts = lo + idx; done = true; break; }
                            idx -= len;
                        }
                    }
                }
            }
            D.row = s_row[warp];
        }
        if (lane != 0) continue;                             // the branchy remainder is one lane's work
        double imm, asym;
        vacc_multipliers(P, c.last_dose_day[ex], c.day.day, imm, asym);
        int32_t rec[MTD_DELTA_WORDS];
#pragma unroll
        for (int k = 0; k < MTD_DELTA_WORDS; k++) rec[k] = 0;
        Traj T;
        sample_trajectory(P, D, c.epi_age_bracket[ex], c.day.day, c.day.day, ts, asym, c.day.testing_enabled != 0,
                          c.day.quarantine_enabled != 0, c.test_day[2 * (int64_t)ex] != MTD_NONE, T);
        const int ns = T.ns, inf_type = T.inf_type, test_day = T.test_day, test_ts = T.test_ts, tres_day = T.tres_day;
        const int start_hosp = T.start_hosp, end_hosp = T.end_hosp;
        const int32_t *sched = T.sched;
        if (D.miss) atomicAdd(&c.ctr->tape_miss, D.miss);
        rec[0] = ex; rec[1] = (int32_t)(uint32_t)key; rec[2] = (int32_t)(uint32_t)(key >> 32);
        rec[3] = inf_type | (ns << 8);
        for (int k = 0; k < ns; k++) rec[4 + k] = sched[k];
        rec[10] = test_day; rec[11] = test_ts; rec[12] = tres_day; rec[13] = test_ts;
        rec[14] = start_hosp; rec[15] = ts; rec[16] = end_hosp;
        int slot = atomicAdd(&dbuf[0], 1);          // record 0 is the header: word 0 = number of records
        if (slot < cap) {
            int4 *dst = (int4 *)(dbuf + (int64_t)(slot + 1) * MTD_DELTA_WORDS);
#pragma unroll
            for (int k = 0; k < MTD_DELTA_WORDS / 4; k++) dst[k] = make_int4(rec[4 * k], rec[4 * k + 1], rec[4 * k + 2], rec[4 * k + 3]);
        } else c.ctr->error = MTD_E_CAPACITY;
    }
}

// K9 -- the itinerary stage's initial cases (itinerary.py:373-395): on the first simulated day every agent that starts
// Infectious gets an exposure day inside the last `pre_days` days and a full trajectory from it
// (Epidemiology.simulate_seir_state_transition with incremental_days < 1, epidemiology.py:233-500).  Transitions dated
// before today are dropped; the agent's states are reset to the states PRECEDING the first transition kept (:243,
// :257-470, as written: an agent still waiting for ExposedToInfectious is (Exposed, PreAsymptomatic) whatever its
// branch, and an agent whose whole trajectory lies in the past stays Exposed with an empty schedule).
// One thread per case; the ids are distinct.  out (optional): 8 words per case
// [seir, infection type, severity, recovered, entries appended, exposure timestep, exposure day, symptomatic day].
__global__ void k9_seed_kernel(DevCtx c, int64_t n, const int32_t *ids, const int32_t *expo_day, int pre_days,
                               const uint8_t *ts_list, int n_ts, const double *asym_mult, const double *draws, int32_t *out,
                               int32_t *status /* [error, schedule overflows, tape misses] */)
{
    const mtd_params *P = c.P;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int ag = ids[i];
        if (ag < 0 || ag >= c.n_agents) { status[0] = MTD_E_ARG; continue; }
        double row[MTD_EXPO_W];
        Draws D; D.miss = 0;
        int ts, incr0;
        if (c.day.mode == MTD_MODE_REPLAY) {
            D.row = draws + i * MTD_EXPO_W;
            ts = isnan(D.row[0]) ? 0 : (int)D.row[0];
            incr0 = expo_day[i];
        } else {
            uint32_t w[4];
            philox4x32((uint32_t)c.day.day, (uint32_t)ag, 0u, 0u, c.day.seed, SITE_EXPO, w);
            ts = ts_list[__umulhi(w[0], (uint32_t)n_ts)];                              // np.random.choice(overlapping_timesteps)
            incr0 = expo_day ? expo_day[i] : c.day.day - pre_days + (int)__umulhi(w[1], (uint32_t)pre_days);
            for (int slot = 1; slot < MTD_EXPO_W; slot++) row[slot] = philox_slot_draw(c, P, ag, slot);
            D.row = row;
        }
        Traj T;
        sample_trajectory(P, D, c.epi_age_bracket[ag], c.day.day, incr0, ts, asym_mult ? asym_mult[i] : 1.0,
                          c.day.testing_enabled != 0, c.day.quarantine_enabled != 0, c.test_day[2 * (int64_t)ag] != MTD_NONE, T);
        if (D.miss) atomicAdd(&status[2], D.miss);
        int cnt = c.sched_cnt[ag];
        int nx = c.sched_next[ag];
        for (int k = 0; k < T.ns; k++) {
            if (cnt < MTD_MAXSCHED) { c.sched[(int64_t)ag * MTD_MAXSCHED + cnt++] = T.sched[k]; nx = min(nx, T.sched[k] >> 16); }
            else atomicAdd(&status[1], 1);
        }
        c.sched_cnt[ag] = (uint8_t)cnt;
        c.sched_next[ag] = (int16_t)nx;
        int seir = S_EXP, it = T.inf_type, sv = SV_UNDEF;
        if (incr0 < 1) {
            switch (T.first_T) {
            case T_EXP_INF:  seir = S_EXP; it = IT_PREASYMP; sv = SV_UNDEF; break;
            case T_INF_SYMP: seir = S_INF; it = IT_PRESYMP; sv = SV_UNDEF; break;
            case T_SYMP_SEV: case T_MILD_REC: seir = S_INF; it = IT_SYMP; sv = SV_MILD; break;
            case T_SEV_CRI:  case T_SEV_REC:  seir = S_INF; it = IT_SYMP; sv = SV_SEVERE; break;
            case T_CRI_DEATH: case T_CRI_REC: seir = S_INF; it = IT_SYMP; sv = SV_CRITICAL; break;
            case T_ASYMP_REC: seir = S_INF; it = IT_ASYMP; sv = SV_UNDEF; break;
            case T_REC_SUS:  seir = S_REC; it = IT_UNDEF; sv = SV_UNDEF; break;
            default: break;
            }
        }
        c.seir[ag] = (int8_t)seir; c.inf_type[ag] = (int8_t)it; c.inf_sev[ag] = (int8_t)sv;
        if (T.test_day != MTD_NONE) {
            c.test_day[2 * (int64_t)ag] = T.test_day; c.test_day[2 * (int64_t)ag + 1] = T.test_ts;
            c.test_result_day[2 * (int64_t)ag] = T.tres_day; c.test_result_day[2 * (int64_t)ag + 1] = T.test_ts;
        }
        if (T.start_hosp != MTD_NONE) {               // schedule_hospitalisation keeps an earlier start day (:651-658)
            int32_t *hp = c.hosp + 3 * (int64_t)ag;
            if (hp[0] == MTD_NONE) hp[0] = T.start_hosp;
            hp[1] = ts; hp[2] = T.end_hosp;
        }
        if (out) {
            int32_t *o = out + 8 * i;
            o[0] = seir; o[1] = it; o[2] = sv; o[3] = T.recovered; o[4] = T.ns; o[5] = ts; o[6] = incr0; o[7] = T.symptomatic_day;
        }
    }
}

// resolution across ranks: earliest key per agent over all gathered records
__global__ void k3_resolve_kernel(DevCtx c, const int32_t *bufs, int n_bufs, int64_t cap)
{
    for (int bi = 0; bi < n_bufs; bi++) {
        const int32_t *buf = bufs + (int64_t)bi * (cap + 1) * MTD_DELTA_WORDS;
        int n = buf[0]; if (n > cap) n = (int)cap;
        for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
            const int32_t *rec = buf + (int64_t)(t + 1) * MTD_DELTA_WORDS;
            unsigned long long key = ((unsigned long long)(uint32_t)rec[2] << 32) | (uint32_t)rec[1];
            atomicMin(&c.win[rec[0]], key);
        }
    }
}

__global__ void k3_apply_kernel(DevCtx c, const int32_t *bufs, int n_bufs, int64_t cap, unsigned long long *expo_list)
{
    for (int bi = 0; bi < n_bufs; bi++) {
        const int32_t *buf = bufs + (int64_t)bi * (cap + 1) * MTD_DELTA_WORDS;
        int n = buf[0]; if (n > cap) n = (int)cap;
        for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
            const int32_t *rec = buf + (int64_t)(t + 1) * MTD_DELTA_WORDS;
            const int ag = rec[0];
            unsigned long long key = ((unsigned long long)(uint32_t)rec[2] << 32) | (uint32_t)rec[1];
            if (c.win[ag] != key) continue;
            int ns = (rec[3] >> 8) & 0xff;
            int cnt = c.sched_cnt[ag];
            int nx = c.sched_next[ag];
            for (int k = 0; k < ns; k++) {            // agent_state_transition_by_day.append(...)
                if (cnt < MTD_MAXSCHED) { c.sched[(int64_t)ag * MTD_MAXSCHED + cnt++] = rec[4 + k]; nx = min(nx, rec[4 + k] >> 16); }
                else atomicAdd(&c.sctr->sched_overflow, 1);
            }
            c.sched_cnt[ag] = (uint8_t)cnt;
            c.sched_next[ag] = (int16_t)nx;
            c.seir[ag] = S_EXP; c.inf_type[ag] = (int8_t)(rec[3] & 0xff); c.inf_sev[ag] = SV_UNDEF;   // :219-225
            if (rec[10] != MTD_NONE) {
                c.test_day[2 * (int64_t)ag] = rec[10]; c.test_day[2 * (int64_t)ag + 1] = rec[11];
                c.test_result_day[2 * (int64_t)ag] = rec[12]; c.test_result_day[2 * (int64_t)ag + 1] = rec[13];
            }
            if (rec[14] != MTD_NONE) {                // schedule_hospitalisation keeps an earlier start day (:651-658)
                int32_t *h = c.hosp + 3 * (int64_t)ag;
                if (h[0] == MTD_NONE) h[0] = rec[14];
                h[1] = rec[15]; h[2] = rec[16];
            }
            int idx = atomicAdd(&c.sctr->n_applied, 1);
            if (idx < c.cap_expo) { expo_list[2 * (int64_t)idx] = key; expo_list[2 * (int64_t)idx + 1] = (unsigned long long)ag; }
        }
    }
}

// The winner table is only meaningful within a day: clear exactly the entries this day touched (the endpoints of the
// successful rows and the agents of every gathered record) instead of refilling all N entries every day.
__global__ void k3_reset_win_kernel(DevCtx c, const int32_t *bufs, int n_bufs, int64_t cap)
{
    const int n = *c.n_succ_pub;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        const RowRec r = c.rows[c.succ[t]];
        c.win[r.a1] = ~0ull; c.win[r.a2] = ~0ull;
    }
    for (int bi = 0; bi < n_bufs; bi++) {
        const int32_t *buf = bufs + (int64_t)bi * (cap + 1) * MTD_DELTA_WORDS;
        int nr = buf[0]; if (nr > cap) nr = (int)cap;
        for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < nr; t += gridDim.x * blockDim.x)
            c.win[buf[(int64_t)(t + 1) * MTD_DELTA_WORDS]] = ~0ull;
    }
}

// seirstateutil.py:105-136
__device__ __forceinline__ void transition(int seir, int it, int sv, int trans, int &ns, int &nt, int &nv)
{
    switch (trans) {
    case T_EXP_INF:  ns = S_INF; nt = (it == IT_PREASYMP) ? IT_ASYMP : it; nv = SV_MILD; return;
    case T_INF_SYMP: ns = IT_SYMP; /* reference quirk B1: InfectionType.Symptomatic (4) returned as the SEIR state */
                     nt = (it == IT_PRESYMP) ? IT_SYMP : it; nv = sv; return;
    case T_SYMP_SEV: ns = seir; nt = it; nv = SV_SEVERE; return;
    case T_SEV_CRI:  ns = seir; nt = it; nv = SV_CRITICAL; return;
    case T_CRI_DEATH: ns = S_DEC; nt = it; nv = sv; return;
    case T_ASYMP_REC: case T_MILD_REC: case T_SEV_REC: case T_CRI_REC: ns = S_REC; nt = IT_UNDEF; nv = SV_UNDEF; return;
    case T_REC_SUS:  ns = S_SUS; nt = it; nv = sv; return;
    default:         ns = S_UNDEF; nt = IT_UNDEF; nv = SV_UNDEF; return;
    }
}

// fire: false = the agent is masked out today (its entry dated today stays, unfired); either way the agent's earliest
// entry day is recomputed over whatever its list holds afterwards
__device__ __forceinline__ void k4_advance_agent(const DevCtx &c, int64_t ag, int day, bool fire)
{
    int32_t *sc = c.sched + ag * MTD_MAXSCHED;
    int cnt = c.sched_cnt[ag];
    int today = -1;
    if (fire) for (int i = 0; i < cnt; i++) if ((sc[i] >> 16) == day) { today = i; break; }   // first entry dated today only
    if (today >= 0) {
        int trans = (sc[today] >> 8) & 0xff;
        int ns, nt, nv;
        transition(c.seir[ag], c.inf_type[ag], c.inf_sev[ag], trans, ns, nt, nv);
        c.seir[ag] = (int8_t)ns; c.inf_type[ag] = (int8_t)nt; c.inf_sev[ag] = (int8_t)nv;
        for (int i = today; i + 1 < cnt; i++) sc[i] = sc[i + 1];                          // del list[today_index]
        cnt--;
        c.sched_cnt[ag] = (uint8_t)cnt;
    }
    // earliest day over what is left: entries of today or earlier that did not fire (a second entry of the same day, a masked
    // agent's entry) keep it at or below today, so the agent is simply looked at again tomorrow
    int nx = MTD_SCHED_NEXT_NONE;
    for (int i = 0; i < cnt; i++) nx = min(nx, sc[i] >> 16);
    c.sched_next[ag] = (int16_t)nx;
}

// K4 reads one 16-bit "earliest entry day" per agent, 8 agents per load: an agent is touched on the day one of its entries
// is due, not on every day it has a pending schedule.  The due agents of a warp's 256 agents are first collected in a
// shared queue and then advanced a lane per agent: at the height of a wave several percent of the population change state
// every day, and advancing them where they are found would leave 31 lanes idle behind each of them.
#define K4_THREADS 256
__global__ void __launch_bounds__(K4_THREADS) k4_advance_kernel(DevCtx c, int day, const uint8_t *mask)
{
    __shared__ int32_t queue[K4_THREADS / 32][256];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int32_t *qw = queue[wid];
    const int64_t n8 = c.n_agents >> 3;
    const uint4 *nx8 = (const uint4 *)c.sched_next;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    // the trip count is the same for every lane of a warp (the loop body holds warp-wide primitives)
    for (int64_t q0 = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31); q0 < n8; q0 += stride) {
        const int64_t q = q0 + lane;
        uint32_t due = 0;                                                     // bit k: agent 8 q + k has an entry due
        if (q < n8) {
            const uint4 v = nx8[q];
            const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const int nx = (int)(int16_t)((w[k >> 1] >> (16 * (k & 1))) & 0xffffu);
                due |= (nx <= day ? 1u : 0u) << k;
            }
        }
        if (__ballot_sync(FULLMASK, due != 0) == 0) continue;
        const int cnt = __popc(due);
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULLMASK, incl, o); if (lane >= o) incl += t; }
        const int total = __shfl_sync(FULLMASK, incl, 31);
        int pos = incl - cnt;
        for (uint32_t b = due; b; b &= b - 1) qw[pos++] = (int32_t)((q << 3) + __ffs(b) - 1);
        __syncwarp();
        for (int i = lane; i < total; i += 32) {
            const int64_t ag = qw[i];
            k4_advance_agent(c, ag, day, !(mask && !mask[ag]));
        }
        __syncwarp();
    }
    if (blockIdx.x == 0 && threadIdx.x < (int)(c.n_agents & 7)) {          // tail
        const int64_t ag = (n8 << 3) + threadIdx.x;
        if ((int)c.sched_next[ag] <= day) k4_advance_agent(c, ag, day, !(mask && !mask[ag]));
    }
}

// sched_next from scratch (after the schedules were uploaded: mtd_set_state, checkpoints)
__global__ void k4_rebuild_next_kernel(DevCtx c)
{
    for (int64_t ag = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; ag < c.n_agents; ag += (int64_t)gridDim.x * blockDim.x) {
        const int cnt = c.sched_cnt[ag];
        int nx = MTD_SCHED_NEXT_NONE;
        for (int i = 0; i < cnt && i < MTD_MAXSCHED; i++) nx = min(nx, c.sched[ag * MTD_MAXSCHED + i] >> 16);
        c.sched_next[ag] = (int16_t)nx;
    }
}

// dynamicstatistics.py:169-193: agents per SEIR state; 16 agents per load, per-thread counters in registers
__global__ void occupancy_kernel(DevCtx c, unsigned long long *out6)
{
    __shared__ unsigned int h[8];
    if (threadIdx.x < 8) h[threadIdx.x] = 0;
    __syncthreads();
    unsigned int cnt[6] = { 0, 0, 0, 0, 0, 0 };
    const int64_t n16 = c.n_agents >> 4;
    const uint4 *s16 = (const uint4 *)c.seir;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n16; q += (int64_t)gridDim.x * blockDim.x) {
        const uint4 v = s16[q];
        const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const uint32_t st = (w[k >> 2] >> (8 * (k & 3))) & 0xffu;
#pragma unroll
            for (int j = 0; j < 6; j++) cnt[j] += (st == (uint32_t)j);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < (int)(c.n_agents & 15)) {
        const int st = c.seir[(n16 << 4) + threadIdx.x];
#pragma unroll
        for (int j = 0; j < 6; j++) cnt[j] += (st == j);
    }
#pragma unroll
    for (int j = 0; j < 6; j++) {
        const unsigned int t = __reduce_add_sync(FULLMASK, cnt[j]);
        if ((threadIdx.x & 31) == 0 && t) atomicAdd(&h[j], t);
    }
    __syncthreads();
    if (threadIdx.x < 6 && h[threadIdx.x]) atomicAdd(&out6[threadIdx.x], (unsigned long long)h[threadIdx.x]);
}

// ------------------------------------------------------------------------------------------------ peer exchange
// Cells are sharded over the GPUs of one box and the per-agent state is replicated (the
// reference's stateful-actor exchange, actor_dist.py:491-518).  Each rank PUSHES the records it produced -- a few
// thousand 80-byte records, exactly as many as there are -- into its slice of every peer's exchange buffer with plain
// stores over NVLink, then raises a per-source arrival flag on the peer; the apply chain of a rank starts once every
// source's flag shows today's sequence number.  Slices are double-buffered by day parity: a rank can only push day
// s+2 after it applied s+1, which needed every peer's push of s+1, which the peer enqueued after its own apply of s.
__global__ void xchg_push_kernel(DevExchange x, int par)
{
    const int64_t slice = (x.cap + 1) * MTD_DELTA_WORDS;                // int32 words per slice
    const int32_t *src = x.buf[x.rank] + ((int64_t)par * x.world + x.rank) * slice;
    int n = src[0]; if (n > x.cap) n = (int)x.cap;
    const int64_t n4 = (int64_t)(n + 1) * (MTD_DELTA_WORDS / 4);
    const int4 *s4 = (const int4 *)src;
    for (int p = 0; p < x.world; p++) {
        if (p == x.rank) continue;
        int4 *d4 = (int4 *)(x.buf[p] + ((int64_t)par * x.world + x.rank) * slice);
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x)
            d4[i] = s4[i];
    }
}

__global__ void xchg_signal_kernel(DevExchange x, int par, unsigned long long seq)
{
    const int p = threadIdx.x;
    if (p >= x.world) return;
    __threadfence_system();
    unsigned long long *f = x.flags[p] + (int64_t)par * x.world + x.rank;
    asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(f), "l"(seq) : "memory");
}

__global__ void xchg_wait_kernel(DevExchange x, int par, unsigned long long seq, StateCounters *sctr, long long timeout_ns)
{
    const int q = threadIdx.x;
    if (q >= x.world) return;
    const unsigned long long *f = x.flags[x.rank] + (int64_t)par * x.world + q;
    long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (true) {
        unsigned long long v;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
        if (v >= seq) break;
        long long t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > timeout_ns) { sctr->error = MTD_E_TIMEOUT; break; }
        __nanosleep(200);
    }
}

// mtd_set_contacts: caller's pairs + overlap lists of ONE cell -> dense rows, one slot
__global__ void set_contacts_kernel(DevCtx c, int cell, int64_t n, const int32_t *a1, const int32_t *a2)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k0 = c.ov_off[i], k1 = c.ov_off[i + 1];
        int dur = 0;
        for (int64_t k = k0; k < k1; k++) dur += (int)c.ov_se[2 * k + 1] - (int)c.ov_se[2 * k] + 1;      // epidemiology.py:170
        RowRec r;
        r.a1 = a1[i]; r.a2 = a2[i]; r.slot = 0;
        const uint32_t st = k1 > k0 ? c.ov_se[2 * k0] : 0u, en = k1 > k0 ? c.ov_se[2 * (k1 - 1) + 1] : 0u;
        r.sed = st | (en << 8) | ((uint32_t)dur << 16);
        c.rows[i] = r;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        c.slot_cell[0] = cell; c.slot_first[0] = 0u; c.slot_m[0] = 0; c.slot_off[0] = 0;
        c.ctr->n_slots = 1; c.ctr->n_rows = (int)n;
    }
}

}  // namespace

int launch_set_contacts(const DevCtx &c, int cell, int64_t n, const int32_t *a1, const int32_t *a2, cudaStream_t st)
{
    set_contacts_kernel<<<64, 256, 0, st>>>(c, cell, n, a1, a2);
    return 1;
}

int launch_k3_eval(const DevCtx &c, int32_t *dbuf, int64_t cap, bool win_clean, cudaStream_t st, int sm_count)
{
    if (!win_clean) cudaMemsetAsync(c.win, 0xff, sizeof(unsigned long long) * (size_t)c.n_agents, st);
    cudaMemsetAsync(dbuf, 0, sizeof(int32_t) * MTD_DELTA_WORDS, st);
    k3_eval_kernel<<<sm_count * 8, 256, 0, st>>>(c);
    k3_trajectory_kernel<<<sm_count * 8, K3T_WARPS * 32, 0, st>>>(c, dbuf, cap);
    return 2;
}

int launch_k3_apply(const DevCtx &c, const int32_t *bufs, int n_bufs, int64_t cap, bool resolve,
                     unsigned long long *expo_list, cudaStream_t st, int sm_count)
{
    if (resolve) k3_resolve_kernel<<<sm_count, 128, 0, st>>>(c, bufs, n_bufs, cap);
    k3_apply_kernel<<<sm_count, 128, 0, st>>>(c, bufs, n_bufs, cap, expo_list);
    k3_reset_win_kernel<<<sm_count, 128, 0, st>>>(c, bufs, n_bufs, cap);
    return resolve ? 3 : 2;
}

int launch_k4_rebuild_next(const DevCtx &c, cudaStream_t st, int sm_count)
{
    k4_rebuild_next_kernel<<<sm_count * 8, 256, 0, st>>>(c);
    return 1;
}

int launch_k4(const DevCtx &c, int32_t day, const uint8_t *mask, cudaStream_t st, int sm_count)
{
    k4_advance_kernel<<<sm_count * 4, K4_THREADS, 0, st>>>(c, day, mask);
    return 1;
}

// push this rank's slice to every peer and raise the arrival flags (main stream, after the trajectory kernel)
int launch_xchg_push(const DevExchange &x, int par, unsigned long long seq, cudaStream_t st)
{
    xchg_push_kernel<<<32, 256, 0, st>>>(x, par);
    xchg_signal_kernel<<<1, 32, 0, st>>>(x, par, seq);
    return 2;
}

// block the state stream until every source's slice of this parity carries sequence number seq
int launch_xchg_wait(const DevExchange &x, int par, unsigned long long seq, StateCounters *sctr, cudaStream_t st)
{
    xchg_wait_kernel<<<1, 32, 0, st>>>(x, par, seq, sctr, 10000000000LL);
    return 1;
}

int launch_occupancy(const DevCtx &c, unsigned long long *out6, cudaStream_t st, int sm_count)
{
    cudaMemsetAsync(out6, 0, 6 * sizeof(unsigned long long), st);
    occupancy_kernel<<<sm_count * 4, 256, 0, st>>>(c, out6);
    return 1;
}

int launch_k9_seed(const DevCtx &c, int64_t n, const int32_t *ids, const int32_t *expo_day, int pre_days, const uint8_t *ts_list,
                   int n_ts, const double *asym_mult, const double *draws, int32_t *out, int32_t *status, cudaStream_t st)
{
    if (n <= 0) return 0;
    const int grid = (int)((n + 127) / 128 < 1184 ? (n + 127) / 128 : 1184);
    k9_seed_kernel<<<grid, 128, 0, st>>>(c, n, ids, expo_day, pre_days, ts_list, n_ts, asym_mult, draws, out, status);
    return 1;
}
