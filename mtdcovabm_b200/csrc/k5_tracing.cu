// K5 -- contact tracing (Epidemiology.contact_tracing, epidemiology.py:667-834).
//   scan      : one thread per contact row of the look-back slices; rows with exactly one traced endpoint draw a keyed
//               uniform against the sim-cell-type success probability -> primary-contact flags (:677-721)
//   order     : replay = the reference's recorded processing order; Philox = ascending agent id (flag compaction)
//   primary   : one thread per primary contact: PositiveContact quarantine + test (:725-735), and the claim
//               "first primary contact of its residence" (atomicMin on the processing index, :750-751)
//   select    : the claiming thread samples round(L * p_residence) co-residents (+ staff) without replacement (:793-795)
//               and claims each for its residence (atomicMin: the reference's global secondary set, :798)
//   secondary : per claimed secondary contact the SecondaryContact quarantine + test draws, applied -- as in the
//               reference -- to the PRIMARY's record (quirk B9, :800-809)
#include "common.cuh"
#include "scan.cuh"

#include "k5_ctx.h"

namespace {

enum { ST_SCAN = 0, ST_MISS = 1, ST_PRIM = 2, ST_SEC = 3 };
enum { Q_POSITIVE_CONTACT = 2, Q_SECONDARY_CONTACT = 3, SITE_TRACE_PRIM = 5 };

__global__ void k5_mark_kernel(TraceCtx c)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < c.n_traced; i += (int64_t)gridDim.x * blockDim.x) {
        int a = c.traced[i];
        if (a >= 0 && a < c.n_agents) c.is_traced[a] = 1;
    }
}

__global__ void k5_scan_kernel(TraceCtx c)
{
    const int64_t n = c.slice_off[c.n_slices];
    int local = 0;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        int back = 0;
        while (back + 1 < c.n_slices && r >= c.slice_off[back + 1]) back++;
        const int sct = c.rows3[3 * r], a1 = c.rows3[3 * r + 1], a2 = c.rows3[3 * r + 2];
        const int f1 = c.is_traced[a1], f2 = c.is_traced[a2];
        if (f1 == f2) continue;                              // exactly one endpoint is a traced positive (:711)
        const uint64_t idx = (uint64_t)(r - c.slice_off[back]);
        double u;
        if (c.day.mode == MTD_MODE_REPLAY) {
            const uint64_t key = ((uint64_t)back << 32) | idx;
            long long lo = 0, hi = c.n_scan;
            while (lo < hi) { long long mid = (lo + hi) >> 1; if (c.scan_key[mid] < key) lo = mid + 1; else hi = mid; }
            if (!(lo < c.n_scan && c.scan_key[lo] == key)) { atomicAdd(&c.stats[ST_MISS], 1); continue; }
            u = c.scan_u[lo];
        } else {
            uint32_t w[4];
            philox4x32((uint32_t)c.day.day, (uint32_t)back, (uint32_t)a1, (uint32_t)a2, c.day.seed, SITE_TRACE, w);   // keyed by the pair,
                                        // not by the row index: the same draw however the contact rows are sharded over ranks
            u = u53(w[0], w[1]);
        }
        local++;
        if (u < c.P->tracing_success_prob[sct]) c.primary[f1 ? a2 : a1] = 1;        // :714-721
    }
    if (local) atomicAdd(&c.stats[ST_SCAN], local);
}

// Philox mode: processing order = ascending agent id
__global__ void k5_compact_kernel(TraceCtx c, const int32_t *excl)
{
    for (int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; a < c.n_agents; a += (int64_t)gridDim.x * blockDim.x) {
        if (a == 0) *c.n_order = excl[c.n_agents];
        if (c.primary[a]) c.order[excl[a]] = (int32_t)a;
    }
}

// replay mode: processing order = the tape's; every recorded primary must be flagged and vice versa
__global__ void k5_order_replay_kernel(TraceCtx c)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < c.n_prim; i += (int64_t)gridDim.x * blockDim.x) {
        if (i == 0) *c.n_order = (int32_t)c.n_prim;
        const int p = c.prim_id[i];
        c.order[i] = p;
        if (p < 0 || p >= c.n_agents || !c.primary[p]) atomicAdd(&c.stats[ST_MISS], 1);
    }
}

struct TraceDraws {
    const TraceCtx *c; int p; const double *rec; int len, pos, miss;
    __device__ double next() { if (pos >= len) { miss++; return 0.0; } return rec[pos++]; }
    __device__ void philox(uint32_t slot, uint32_t w[4]) const
    {
        philox4x32((uint32_t)c->day.day, (uint32_t)p, slot, 0u, c->day.seed, SITE_TRACE_PRIM, w);
    }
    __device__ int ts(uint32_t slot)
    {
        if (c->day.mode == MTD_MODE_REPLAY) return (int)next();
        uint32_t w[4]; philox(slot, w);
        return 42 + (int)__umulhi(w[0], 79u);
    }
    __device__ double uniform(uint32_t slot)
    {
        if (c->day.mode == MTD_MODE_REPLAY) return next();
        uint32_t w[4]; philox(slot, w);
        return u53(w[0], w[1]);
    }
    __device__ int lognormal_days(uint32_t slot, double mu, double sigma)
    {
        if (c->day.mode == MTD_MODE_REPLAY) return (int)next();
        uint32_t w[4]; philox(slot, w);
        double u1 = u53(w[0], w[1]), u2 = u53(w[2], w[3]);
        double z = sqrt(-2.0 * log(1.0 - u1)) * cos(6.283185307179586 * u2);
        return (int)rint(exp(mu + sigma * z));
    }
};

// schedule_quarantine (epidemiology.py:568-625) for PositiveContact / SecondaryContact on record r
__device__ void trace_quarantine(TraceDraws &T, int r, int ts, int qtype, uint32_t uslot)
{
    const TraceCtx &c = *T.c;
    if (!c.day.quarantine_enabled) return;
    const double u = T.uniform(uslot);
    if (u < c.P->ignore_quarantine_rules_probability) return;
    int start = c.day.day, end;
    if (qtype == Q_POSITIVE_CONTACT) { start += c.P->contact_tracing_positive_delay_days; end = start + c.P->quarantine_positive_contact_duration; }
    else { start += c.P->contact_tracing_secondary_delay_days; end = start + c.P->quarantine_secondary_contact_duration; }
    int32_t *q = c.quarantine + 3 * (int64_t)r;
    if (q[0] != MTD_NONE && start >= q[0]) return;              // :606-611
    q[0] = start; q[1] = ts; q[2] = end;
}

// schedule_test (epidemiology.py:502-565)
__device__ void trace_test(TraceDraws &T, int r, int ts, int qtype, uint32_t lnslot)
{
    const TraceCtx &c = *T.c;
    if (!c.day.testing_enabled) return;
    int32_t *td = c.test_day + 2 * (int64_t)r, *tr = c.test_result_day + 2 * (int64_t)r;
    if (td[0] != MTD_NONE) return;                              // :519-521
    int days = (qtype == Q_POSITIVE_CONTACT) ? c.P->contact_tracing_positive_delay_days : c.P->contact_tracing_secondary_delay_days;
    days += T.lognormal_days(lnslot, c.P->test_mu, c.P->test_sigma);
    const int tday = c.day.day + days;
    const int rday = tday + T.lognormal_days(lnslot + 1, c.P->result_mu, c.P->result_sigma);
    td[0] = tday; td[1] = ts; tr[0] = rday; tr[1] = ts;
}

__device__ __forceinline__ TraceDraws open_draws(const TraceCtx &c, int i, int p)
{
    TraceDraws T; T.c = &c; T.p = p; T.rec = nullptr; T.len = 0; T.pos = 0; T.miss = 0;
    if (c.day.mode == MTD_MODE_REPLAY && i < c.n_prim) {
        T.rec = c.prim_draws + c.prim_off[i]; T.len = c.prim_off[i + 1] - c.prim_off[i]; T.pos = c.cursor[i];
    }
    return T;
}

__device__ __forceinline__ bool valid_res(const TraceCtx &c, int res)
{
    return res >= 0 && res < c.n_res_cells && c.res_off[res + 1] > c.res_off[res];
}

__global__ void k5_primary_kernel(TraceCtx c)
{
    const int n = *c.n_order;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int p = c.order[i];
        if (p < 0 || p >= c.n_agents) continue;
        if (c.epi_present && !c.epi_present[p]) continue;           // agents_epi[contact_id] missing (:727-733)
        TraceDraws T = open_draws(c, i, p);
        int ts = T.ts(0);
        trace_quarantine(T, p, ts, Q_POSITIVE_CONTACT, 1);
        ts = T.ts(2);
        trace_test(T, p, ts, Q_POSITIVE_CONTACT, 3);
        c.cursor[i] = T.pos;
        const int res = c.res_cellid[p];
        if (valid_res(c, res)) atomicMin(&c.res_first[res], i);     // res_cell_ids: first primary of the residence wins
        else if (c.day.mode == MTD_MODE_REPLAY && T.pos != T.len) T.miss++;
        if (T.miss) atomicAdd(&c.stats[ST_MISS], T.miss);
    }
}

__global__ void k5_select_kernel(TraceCtx c)
{
    const int n = *c.n_order;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int p = c.order[i];
        if (p < 0 || p >= c.n_agents) continue;
        if (c.epi_present && !c.epi_present[p]) continue;
        const int res = c.res_cellid[p];
        if (!valid_res(c, res)) continue;
        TraceDraws T = open_draws(c, i, p);
        if (c.res_first[res] != i) {                                // another primary of this residence came first
            if (c.day.mode == MTD_MODE_REPLAY && T.pos != T.len) atomicAdd(&c.stats[ST_MISS], 1);
            continue;
        }
        // sec_contact_ids = residents without the contact itself, then staff (:767-788)
        const int64_t b = c.res_off[res], e = c.res_off[res + 1];
        int32_t *sel = c.sel + b;
        int L = 0;
        for (int64_t q = b; q < e; q++) {
            int m = c.res_members[q];
            if (q - b < c.res_nres[res] && m == p) continue;
            sel[L++] = m;
        }
        int k = 0;
        if (L > 0) {
            k = (int)rint((double)L * c.P->tracing_success_prob[0]);            // round(len * p_residence) (:793)
            if (k > L) k = L;
            for (int t = 0; t < k; t++) {                                       // np.random.choice(.., replace=False)
                int r = t;
                if (c.day.mode == MTD_MODE_REPLAY) {
                    const int id = (int)T.next();
                    while (r < L && sel[r] != id) r++;
                    if (r == L) { T.miss++; r = t; }
                } else {
                    uint32_t w[4]; T.philox(16u + (uint32_t)t, w);
                    r = t + (int)__umulhi(w[0], (uint32_t)(L - t));
                }
                const int y = sel[r];
                for (int j = r; j > t; j--) sel[j] = sel[j - 1];
                sel[t] = y;
            }
            for (int j = 0; j < k; j++) {
                const int s = sel[j];
                if (s >= 0 && s < c.n_agents && !c.primary[s]) atomicMin(&c.sec_owner[s], i);
            }
        }
        c.res_k[res] = k;
        c.cursor[i] = T.pos;
        if (T.miss) atomicAdd(&c.stats[ST_MISS], T.miss);
    }
}

__global__ void k5_secondary_kernel(TraceCtx c)
{
    const int n = *c.n_order;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int p = c.order[i];
        if (p < 0 || p >= c.n_agents) continue;
        if (c.epi_present && !c.epi_present[p]) continue;
        const int res = c.res_cellid[p];
        if (!valid_res(c, res) || c.res_first[res] != i) continue;
        TraceDraws T = open_draws(c, i, p);
        const int32_t *sel = c.sel + c.res_off[res];
        const int k = c.res_k[res];
        for (int j = 0; j < k; j++) {
            const int s = sel[j];
            if (s < 0 || s >= c.n_agents || c.primary[s] || c.sec_owner[s] != i || c.secondary[s]) continue;   // :798
            // quirk B9: the PRIMARY's record is offered to schedule_quarantine / schedule_test (:800-809)
            const uint32_t base = 65536u + 8u * (uint32_t)j;
            const int t1 = T.ts(base);
            trace_quarantine(T, p, t1, Q_SECONDARY_CONTACT, base + 1);
            const int t2 = T.ts(base + 2);
            trace_test(T, p, t2, Q_SECONDARY_CONTACT, base + 3);
            c.secondary[s] = 1;
        }
        if (c.day.mode == MTD_MODE_REPLAY && T.pos != T.len) T.miss++;
        if (T.miss) atomicAdd(&c.stats[ST_MISS], T.miss);
    }
}

__global__ void k5_count_kernel(TraceCtx c)
{
    int np = 0, ns = 0;
    for (int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; a < c.n_agents; a += (int64_t)gridDim.x * blockDim.x) {
        np += c.primary[a]; ns += c.secondary[a];
    }
    for (int o = 16; o > 0; o >>= 1) { np += __shfl_xor_sync(FULLMASK, np, o); ns += __shfl_xor_sync(FULLMASK, ns, o); }
    if ((threadIdx.x & 31) == 0 && (np | ns)) { atomicAdd(&c.stats[ST_PRIM], np); atomicAdd(&c.stats[ST_SEC], ns); }
}

// K6 -- intervention bookkeeping of the itinerary stage, one thread per agent (Itinerary.sample_intervention_activities,
// itinerary.py:2283-2528), in the reference's order:
//   (A) :2301-2332  a test result due today flags the agent positive unless a false negative (Exposed / Infectious) or
//       only on a false positive (everybody else); a positive agent joins contact_tracing_agent_ids and gets a Positive
//       quarantine (schedule_quarantine, epidemiology.py:568-625); a result day in the past is cleared;
//   (B) :2334-2397  hospitalisation_days are cleared on their end day (not on a day that is also their start day);
//   (C) :2399-2457  quarantine_days are cleared on their end day, unless the window starts today or the hospitalisation
//       ended today inside the window;
//   (D) :2459-2487  test_day is cleared six days after the test.
// Draws: replay tape keyed by agent (two doubles: result draw, quarantine draw) or Philox keyed (day, agent), site 6.
__global__ void k6_test_results_kernel(DevCtx c, const uint8_t *epi_present, uint8_t *positive, int64_t rp_n,
                                       const int32_t *rp_agent, const double *rp_u)
{
    const mtd_params *P = c.P;
    const int day = c.day.day;
    for (int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; a < c.n_agents; a += (int64_t)gridDim.x * blockDim.x) {
        positive[a] = 0;
        if (epi_present && !epi_present[a]) continue;
        int32_t *tr = c.test_result_day + 2 * a;
        const int trd = tr[0];
        if (trd != MTD_NONE) {                                              // (A)
            if (day == trd) {
                const int ts = tr[1];
                double u, u2;
                if (c.day.mode == MTD_MODE_REPLAY) {
                    int64_t lo = 0, hi = rp_n;
                    while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (rp_agent[mid] < (int32_t)a) lo = mid + 1; else hi = mid; }
                    if (lo < rp_n && rp_agent[lo] == (int32_t)a) { u = rp_u[2 * lo]; u2 = rp_u[2 * lo + 1]; }
                    else { u = 2.0; u2 = 2.0; atomicAdd(&c.ctr->tape_miss, 1); }
                } else {
                    uint32_t w[4];
                    philox4x32((uint32_t)day, (uint32_t)a, 0u, 0u, c.day.seed, 6u, w);
                    u = u53(w[0], w[1]); u2 = u53(w[2], w[3]);
                }
                const int st = c.seir[a];
                const bool pos = (st == S_EXP || st == S_INF) ? !(u < P->testing_false_negative_rate)
                                                               : (u < P->testing_false_positive_rate);
                if (pos) {
                    positive[a] = 1;
                    if (c.day.quarantine_enabled && !(u2 < P->ignore_quarantine_rules_probability)) {
                        int32_t *q = c.quarantine + 3 * a;
                        const int start = day, end = day + P->quarantine_positive_duration;
                        if (!(q[0] != MTD_NONE && start >= q[0])) { q[0] = start; q[1] = ts; q[2] = end; }   // epidemiology.py:606-611
                    }
                }
            } else if (day > trd) { tr[0] = MTD_NONE; tr[1] = MTD_NONE; }
        }
        bool hosp_end_today = false;
        int32_t *hp = c.hosp + 3 * a;
        const int h0 = hp[0];
        if (h0 != MTD_NONE && h0 != day && hp[2] == day) { hosp_end_today = true; hp[0] = MTD_NONE; hp[1] = MTD_NONE; hp[2] = MTD_NONE; }   // (B)
        int32_t *q = c.quarantine + 3 * a;
        const int q0 = q[0];
        if (q0 != MTD_NONE) {                                               // (C)
            const int q2 = q[2];
            if (!(q0 == day || (hosp_end_today && day <= q2)) && q2 == day) { q[0] = MTD_NONE; q[1] = MTD_NONE; q[2] = MTD_NONE; }
        }
        int32_t *td = c.test_day + 2 * a;
        const int t0 = td[0];
        if (t0 != MTD_NONE && day != t0 && day - t0 == 6) { td[0] = MTD_NONE; td[1] = MTD_NONE; }   // (D)
    }
}

// flagged agents -> ascending id list (flag scan + compaction), count to n_out
__global__ void k6_compact_kernel(int64_t n_agents, const uint8_t *positive, const int32_t *excl, int32_t *ids, int64_t cap,
                                  int32_t *n_out)
{
    for (int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; a < n_agents; a += (int64_t)gridDim.x * blockDim.x)
        if (positive[a] && excl[a] < cap) ids[excl[a]] = (int32_t)a;
    if (blockIdx.x == 0 && threadIdx.x == 0) *n_out = excl[n_agents];
}

}  // namespace

int launch_k6(const DevCtx &c, const uint8_t *epi_present, uint8_t *positive, int32_t *flag_scan, long long *scan_tmp,
              int32_t *ids, int64_t cap, int32_t *n_out, int64_t rp_n, const int32_t *rp_agent, const double *rp_u,
              cudaStream_t st, int sm_count)
{
    k6_test_results_kernel<<<sm_count * 8, 256, 0, st>>>(c, epi_present, positive, rp_n, rp_agent, rp_u);
    int n = 1 + device_scan<uint8_t, int32_t>(positive, flag_scan, c.n_agents, nullptr, scan_tmp, st);
    k6_compact_kernel<<<sm_count * 8, 256, 0, st>>>(c.n_agents, positive, flag_scan, ids, cap, n_out);
    return n + 1;
}

// phase 1: traced marks + scan of the look-back slices -> primary-contact flags (c.primary).  Under cell sharding every rank
// scans the rows of its own cells; the flags are OR-ed over the ranks before phase 2.
int launch_k5_scan(const TraceCtx &c, cudaStream_t st, int sm_count)
{
    const int T = 256;
    const size_t N = (size_t)c.n_agents;
    cudaMemsetAsync(c.is_traced, 0, N, st); cudaMemsetAsync(c.primary, 0, N, st); cudaMemsetAsync(c.secondary, 0, N, st);
    cudaMemsetAsync(c.stats, 0, 4 * sizeof(int32_t), st); cudaMemsetAsync(c.n_order, 0, sizeof(int32_t), st);
    auto grid = [&](int64_t n) { int64_t g = (n + T - 1) / T; int64_t cap = (int64_t)sm_count * 16; return (unsigned)(g < 1 ? 1 : (g < cap ? g : cap)); };
    int launches = 0;
    if (c.n_traced > 0) { k5_mark_kernel<<<grid(c.n_traced), T, 0, st>>>(c); launches++; }
    const int64_t n_rows = c.slice_off[c.n_slices];
    if (n_rows > 0) { k5_scan_kernel<<<grid(n_rows), T, 0, st>>>(c); launches++; }
    return launches;
}

// phase 2: processing order, primary / secondary scheduling on the (replicated) per-agent records
int launch_k5_finish(const TraceCtx &c, int32_t *flag_scan, long long *scan_tmp, cudaStream_t st, int sm_count)
{
    const int T = 256;
    const size_t N = (size_t)c.n_agents;
    cudaMemsetAsync(c.cursor, 0, sizeof(int32_t) * N, st);
    cudaMemsetAsync(c.sec_owner, 0x7f, sizeof(int32_t) * N, st);
    cudaMemsetAsync(c.res_first, 0x7f, sizeof(int32_t) * (size_t)c.n_res_cells, st);
    cudaMemsetAsync(c.res_k, 0, sizeof(int32_t) * (size_t)c.n_res_cells, st);
    auto grid = [&](int64_t n) { int64_t g = (n + T - 1) / T; int64_t cap = (int64_t)sm_count * 16; return (unsigned)(g < 1 ? 1 : (g < cap ? g : cap)); };
    int launches = 0;
    if (c.day.mode == MTD_MODE_REPLAY) {
        if (c.n_prim > 0) { k5_order_replay_kernel<<<grid(c.n_prim), T, 0, st>>>(c); launches++; }
    } else {
        launches += device_scan<uint8_t, int32_t>(c.primary, flag_scan, c.n_agents, nullptr, scan_tmp, st);
        k5_compact_kernel<<<grid(c.n_agents), T, 0, st>>>(c, flag_scan); launches++;
    }
    // the three phases run over at most n_agents primaries; the kernels read the true count from the device
    const unsigned g = (unsigned)(sm_count * 4);
    k5_primary_kernel<<<g, T, 0, st>>>(c);
    k5_select_kernel<<<g, T, 0, st>>>(c);
    k5_secondary_kernel<<<g, T, 0, st>>>(c);
    k5_count_kernel<<<grid(c.n_agents), T, 0, st>>>(c);
    return launches + 4;
}

int launch_k5(const TraceCtx &c, int32_t *flag_scan, long long *scan_tmp, cudaStream_t st, int sm_count)
{
    return launch_k5_scan(c, st, sm_count) + launch_k5_finish(c, flag_scan, scan_tmp, st, sm_count);
}
