// K1, LEGACY path for cells of 33..1024 intervals: one WARP per cell (round 1's kernel, verified on B200 against the
// reference's golden days).  Kept selectable (MTD_K1_IMPL=legacy) beside the block-per-cell kernels of k1_block.cu, which
// replace it: same inputs, same outputs, bit for bit.  See k1_block.cu for the algorithm description; the differences are
// the adjacency build (pairwise mask AND + ballot here, O(m^2) word operations; start / end timestep tables there), where
// the bit matrix lives (an L2-resident global scratch here, shared memory there) and who runs the non-sequential phases
// (the one warp here, all warps of the block there).
#include "common.cuh"
#include <stdlib.h>

namespace {

__host__ __device__ constexpr int pow2_ge(int x) { int p = 1; while (p < x) p <<= 1; return p; }
__host__ __device__ constexpr size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

// position of the (r+1)-th set bit of v (r is 0-based, popc(v) > r)
__device__ __forceinline__ int nth_set_bit(uint32_t v, int r)
{
    int pos = 0, c;
    c = __popc(v & 0xffffu); if (r >= c) { r -= c; pos += 16; v >>= 16; }
    c = __popc(v & 0xffu);   if (r >= c) { r -= c; pos += 8;  v >>= 8; }
    c = __popc(v & 0xfu);    if (r >= c) { r -= c; pos += 4;  v >>= 4; }
    c = __popc(v & 0x3u);    if (r >= c) { r -= c; pos += 2;  v >>= 2; }
    c = (int)(v & 1u);       if (r >= c) { pos += 1; }
    return pos;
}

// word k of the occupancy mask of the inclusive timestep range [s, e] (timesteps 0..159, five 32-bit words)
__device__ __forceinline__ uint32_t range_word(int s, int e, int k)
{
    int lo = max(s - 32 * k, 0), hi = min(e - 32 * k, 31);
    if (lo > hi) return 0u;
    return (0xffffffffu >> (31 - (hi - lo))) << lo;
}

__device__ __forceinline__ long long tape_find_sel(const DevTape &T, int cell, int agent)
{
    uint64_t key = ((uint64_t)(uint32_t)cell << 32) | (uint32_t)agent;
    long long lo = 0, hi = T.n_sel;
    while (lo < hi) { long long mid = (lo + hi) >> 1; if (T.sel_key[mid] < key) lo = mid + 1; else hi = mid; }
    return (lo < T.n_sel && T.sel_key[lo] == key) ? lo : -1;
}

__device__ __forceinline__ double log_int(const DevCtx &g, int v)
{   // math.log of a small integer: host (glibc) table, the same libm CPython calls
    return (v >= 0 && v < MTD_LOGTAB_N) ? g.logtab[v] : log((double)v);
}

// one 16-byte gather per agent: {soc_rate, age_bracket}
__device__ __forceinline__ AgentAttr load_attr(const DevCtx &g, int ag)
{
    const int4 v = __ldg((const int4 *)(g.k1_attr + ag));
    AgentAttr a;
    a.soc_rate = __hiloint2double(v.y, v.x); a.age_bracket = v.z; a.pad = 0;
    return a;
}

// contactnetwork.py:225-252: float degree of one agent
__device__ __forceinline__ double degree_f2(const DevCtx &g, const mtd_params *P, int ab, double soc, int act, int ts, int pc,
                                            double avg_ts, double avg_pc, double log_avg_ts, double log_avg_pc,
                                            double one_minus_mhd)
{
    if (ab < 0) ab += 16;                                   // numpy negative index
    double cmv = P->cm[ab][act];
    double tsm = 1.0, pcm = 1.0;
    if ((double)ts != avg_ts && avg_ts > 1) tsm = log_int(g, ts) / log_avg_ts;       // math.log(x, base)
    if ((double)pc != avg_pc && avg_pc > 1) pcm = log_int(g, pc) / log_avg_pc;
    return cmv * tsm * pcm * soc * one_minus_mhd;
}

__device__ __forceinline__ double degree_f(const DevCtx &g, const mtd_params *P, int ag, int act, int ts, int pc,
                                           double avg_ts, double avg_pc, double log_avg_ts, double log_avg_pc,
                                           double one_minus_mhd)
{
    const AgentAttr at = load_attr(g, ag);
    return degree_f2(g, P, at.age_bracket, at.soc_rate, act, ts, pc, avg_ts, avg_pc, log_avg_ts, log_avg_pc, one_minus_mhd);
}

__device__ __forceinline__ int degree_i(double d, long long deg_sum, long long max_degree, int pc)
{   // contactnetwork.py:254-269 (Python round = rint)
    long long r;
    if (deg_sum > max_degree) r = (long long)rint((d / (double)deg_sum) * (double)max_degree);
    else r = (long long)rint(d);
    return (int)(r > pc ? pc : r);
}


// ------------------------------------------------------------------------------------------------ large cells
// One warp per cell.  Shared memory per warp is kept small (about 14.5 KB for 320 intervals) so that many cells are in
// flight per SM: the scratch region U is reused phase by phase (sort keys + agent hash -> occupancy masks -> float
// degrees -> greedy order keys), and the adjacency bit matrix -- written once, each row read once when its agent
// becomes the main agent -- lives in a per-warp slice of an L2-resident global scratch, word-major ([word][agent], so
// the writes are coalesced), with the next main agent's row prefetched one iteration ahead.
template <int MM>
struct Lay {
    static constexpr int P2 = pow2_ge(MM);
    static constexpr int BW = (MM + 31) / 32;              // bit-matrix words per row
    static constexpr int H = 2 * P2;                       // hash slots
    static constexpr size_t u_sort = sizeof(uint64_t) * P2 + sizeof(uint32_t) * H;   // keys (or 2 x u32[P2]) + hash
    static constexpr size_t u_mask = sizeof(uint32_t) * 6 * MM;                       // u32[MM][6] occupancy masks
    static constexpr size_t u_bytes = u_sort > u_mask ? u_sort : u_mask;
    static constexpr size_t o_u = 0;
    static constexpr size_t o_hash = sizeof(uint64_t) * P2;                         // inside U
    static constexpr size_t o_agent = o_u + align16(u_bytes);                       // i32[MM] agent id per interval,
                                                                                    //   then per local agent
    static constexpr size_t o_se = o_agent + align16(sizeof(int32_t) * MM);         // u16[MM] start | end << 8
    static constexpr size_t o_la = o_se + align16(2 * MM);                          // u16[MM] local agent of interval,
                                                                                    //   then the greedy order
    static constexpr size_t o_apos = o_la + align16(2 * MM);                        // u16[MM] interval lists
    static constexpr size_t o_aoff = o_apos + align16(2 * MM);                      // u16[MM+1]
    static constexpr size_t o_first = o_aoff + align16(2 * (MM + 1));               // u16[MM] first interval of agent
    static constexpr size_t o_pc = o_first + align16(2 * MM);                       // u16[MM] potential contacts,
                                                                                    //   then the replay picks
    static constexpr size_t o_deg = o_pc + align16(2 * MM);                         // i16[MM]
    static constexpr size_t o_ts = o_deg + align16(2 * MM);                         // u16[MM] timesteps, then draws
    static constexpr size_t o_alive = o_ts + align16(2 * MM);                       // u32[32] agents with degree > 0
    static constexpr size_t o_candw = o_alive + 128;                                // u32[32] candidates of the main
    static constexpr size_t o_ccnt = o_candw + 128;                                 // u32[32] popcounts of candw
    static constexpr size_t bytes = o_ccnt + 128;
    static constexpr size_t bits_words = (size_t)MM * BW;                           // global scratch per warp
};

struct Cell {
    uint64_t *key; int32_t *agent; uint16_t *se, *la, *apos, *aoff, *first, *pc; int16_t *deg; uint16_t *ts;
    uint32_t *mask, *hash, *alive, *candw, *ccnt;
};

__device__ __forceinline__ void warp_bitonic_sort(uint64_t *k, int P2, int lane)
{
    for (int size = 2; size <= P2; size <<= 1) {
        for (int j = size >> 1; j > 0; j >>= 1) {
            for (int t = lane; t < (P2 >> 1); t += 32) {
                int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                int p = i | j;
                bool asc = ((i & size) == 0);
                uint64_t a = k[i], b = k[p];
                if ((a > b) == asc) { k[i] = b; k[p] = a; }
            }
            __syncwarp();
        }
    }
}

// numpy float64 add.reduce (pairwise summation, 8 accumulators, blocks of <= 128), executed by one lane.
__device__ double np_pairwise_sum(const double *a, int n)
{
    if (n < 8) {
        double r = -0.0;
        for (int i = 0; i < n; i++) r += a[i];
        return r;
    }
    if (n <= 128) {
        double r0 = a[0], r1 = a[1], r2 = a[2], r3 = a[3], r4 = a[4], r5 = a[5], r6 = a[6], r7 = a[7];
        int i;
        for (i = 8; i < n - (n % 8); i += 8) {
            r0 += a[i]; r1 += a[i + 1]; r2 += a[i + 2]; r3 += a[i + 3];
            r4 += a[i + 4]; r5 += a[i + 5]; r6 += a[i + 6]; r7 += a[i + 7];
        }
        double res = ((r0 + r1) + (r2 + r3)) + ((r4 + r5) + (r6 + r7));
        for (; i < n; i++) res += a[i];
        return res;
    }
    int n2 = n / 2;
    n2 -= n2 % 8;
    return np_pairwise_sum(a, n2) + np_pairwise_sum(a + n2, n - n2);
}

// Contact row of the agents with local indices x, y (interval lists aoff/apos in position order), in ONE pass over the
// interval pairs.  The reference's key orientation is decided by whose interval is reached first by the ordered double
// loop (contactnetwork.py:160-196): orientation A = x outer / y inner is this loop's own order; orientation B = y outer /
// x inner is the lexicographic order of (pb, pa).  start = first overlap's start, end = last overlap's end, in the
// chosen orientation (contactnetwork.py:101); dur = summed overlap length (epidemiology.py:170).
__device__ __forceinline__ void pair_row_lists(const Cell &c, int x, int y, bool &x_first, int &start, int &end, int &dur)
{
    int fx = 1 << 30, fy = 1 << 30;
    int a_start = 0, a_end = 0; bool a_none = true;
    int b_min = 1 << 30, b_max = -1, b_start = 0, b_end = 0;
    dur = 0;
    const int xa = c.aoff[x], xb = c.aoff[x + 1], ya = c.aoff[y], yb = c.aoff[y + 1];
    if (xb - xa == 1 && yb - ya == 1) {                             // one interval each (the common case): one overlap
        const int pa = c.apos[xa], pb = c.apos[ya];
        start = max((int)(c.se[pa] & 0xff), (int)(c.se[pb] & 0xff)); end = min((int)(c.se[pa] >> 8), (int)(c.se[pb] >> 8));
        dur = end - start + 1; x_first = pa < pb;
        return;
    }
    for (int qa = xa; qa < xb; qa++) {
        const int pa = c.apos[qa]; const int sa = c.se[pa] & 0xff, ea = c.se[pa] >> 8;
        for (int qb = ya; qb < yb; qb++) {
            const int pb = c.apos[qb]; const int sb = c.se[pb] & 0xff, eb = c.se[pb] >> 8;
            const int lo = max(sa, sb), hi = min(ea, eb);
            if (lo <= hi) {
                fx = min(fx, pa); fy = min(fy, pb); dur += hi - lo + 1;
                if (a_none) { a_start = lo; a_none = false; }
                a_end = hi;
                const int kb = (pb << 12) | pa;
                if (kb < b_min) { b_min = kb; b_start = lo; }
                if (kb > b_max) { b_max = kb; b_end = hi; }
            }
        }
    }
    x_first = fx < fy;
    start = x_first ? a_start : b_start;
    end = x_first ? a_end : b_end;
}

template <int MM>
__device__ void process_cell_large(const DevCtx &g, unsigned char *sm, uint32_t *bits, int slot, int lane,
                                   unsigned long long &acc_pairs, unsigned long long &acc_pot)
{
    using L = Lay<MM>;
    Cell c;
    c.key = (uint64_t *)(sm + L::o_u); c.agent = (int32_t *)(sm + L::o_agent); c.se = (uint16_t *)(sm + L::o_se);
    c.la = (uint16_t *)(sm + L::o_la); c.apos = (uint16_t *)(sm + L::o_apos); c.aoff = (uint16_t *)(sm + L::o_aoff);
    c.first = (uint16_t *)(sm + L::o_first); c.pc = (uint16_t *)(sm + L::o_pc); c.deg = (int16_t *)(sm + L::o_deg);
    c.ts = (uint16_t *)(sm + L::o_ts); c.mask = (uint32_t *)(sm + L::o_u); c.hash = (uint32_t *)(sm + L::o_hash);
    c.alive = (uint32_t *)(sm + L::o_alive); c.candw = (uint32_t *)(sm + L::o_candw); c.ccnt = (uint32_t *)(sm + L::o_ccnt);

    const int m = g.slot_m[slot];
    const int64_t base = g.slot_off[slot];
    const int cell = g.slot_cell[slot];
    const int act = g.slot_act[slot];
    uint32_t *t32a = (uint32_t *)c.key;           // two u32[P2] temporaries aliasing the sort-key region
    uint32_t *t32b = t32a + L::P2;
    double *degf = (double *)c.key;

    // ---- 1. restore list (emission) order inside the cell: sort by emission index ----
    int P2 = 2;
    while (P2 < m) P2 <<= 1;
    bool sorted = true;
    for (int p = lane; p < P2; p += 32) {
        uint64_t k = ~0ull;
        if (p < m) {
            uint32_t o = g.csr[base + p].order;
            k = ((uint64_t)o << 32) | (uint32_t)p;
            if (p > 0 && g.csr[base + p - 1].order > o) sorted = false;
        }
        c.key[p] = k;
    }
    sorted = __all_sync(FULLMASK, sorted);
    __syncwarp();
    if (!sorted) warp_bitonic_sort(c.key, P2, lane);
    for (int p = lane; p < m; p += 32) {
        IvRec r = g.csr[base + (uint32_t)c.key[p]];
        c.agent[p] = r.agent; c.se[p] = (uint16_t)(r.se & 0xffffu);
    }
    __syncwarp();

    // ---- 2. local agent index = order of first appearance (contactnetwork.py:161-163); per-agent interval lists ----
    const int H = 2 * P2;
    for (int h = lane; h < H; h += 32) c.hash[h] = 0;
    for (int p = lane; p < P2; p += 32) { t32a[p] = 0xffffffffu; t32b[p] = 0; }
    __syncwarp();
    for (int p = lane; p < m; p += 32) {                 // representative position of every agent
        int ag = c.agent[p];
        uint32_t h = ((uint32_t)ag * 2654435761u) & (H - 1);
        int rep;
        while (true) {
            uint32_t v = atomicCAS(&c.hash[h], 0u, (uint32_t)p + 1u);
            if (v == 0) { rep = p; break; }
            if (c.agent[v - 1] == ag) { rep = (int)v - 1; break; }
            h = (h + 1) & (H - 1);
        }
        c.la[p] = (uint16_t)rep;
    }
    __syncwarp();
    for (int p = lane; p < m; p += 32) atomicMin(&t32a[c.la[p]], (uint32_t)p);      // first position per representative
    __syncwarp();
    int n = 0;
    for (int p0 = 0; p0 < m; p0 += 32) {                 // number first occurrences in position order
        int p = p0 + lane;
        bool isfirst = (p < m) && (t32a[c.la[p]] == (uint32_t)p);
        uint32_t bm = __ballot_sync(FULLMASK, isfirst);
        if (isfirst) { int l = n + __popc(bm & lanemask_lt()); c.first[l] = (uint16_t)p; c.pc[p] = (uint16_t)l; }
        n += __popc(bm);
    }
    __syncwarp();
    for (int p0 = 0; p0 < m; p0 += 32) {
        int p = p0 + lane;
        int l = 0;
        if (p < m) l = c.pc[t32a[c.la[p]]];              // c.pc temporarily maps first position -> local index
        __syncwarp();
        if (p < m) { c.la[p] = (uint16_t)l; atomicAdd(&t32b[l], 1u); }
    }
    __syncwarp();
    {   // exclusive scan of the per-agent interval counts -> aoff; cursor in t32a
        int run = 0;
        unsigned long long sq = 0;
        for (int l0 = 0; l0 < n; l0 += 32) {
            int l = l0 + lane;
            int v = (l < n) ? (int)t32b[l] : 0;
            sq += (unsigned long long)v * v;
            int incl = v;
            for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(FULLMASK, incl, o); if (lane >= o) incl += t; }
            if (l < n) { c.aoff[l] = (uint16_t)(run + incl - v); t32a[l] = run + incl - v; }
            run += __shfl_sync(FULLMASK, incl, 31);
        }
        if (lane == 0) c.aoff[n] = (uint16_t)run;
        for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(FULLMASK, sq, o);
        acc_pairs += ((unsigned long long)m * m - sq) / 2;
        if (lane == 0) atomicAdd(&g.ctr->n_incidences, (unsigned long long)n);
    }
    __syncwarp();
    for (int p0 = 0; p0 < m; p0 += 32) {                 // fill the per-agent lists in position order
        int p = p0 + lane;
        int l = (p < m) ? (int)c.la[p] : (0x10000 + lane);
        uint32_t mmm = __match_any_sync(FULLMASK, l);
        if (p < m) {
            int rank = __popc(mmm & lanemask_lt());
            c.apos[t32a[l] + rank] = (uint16_t)p;
        }
        __syncwarp();
        if (p < m && (__ffs(mmm) - 1) == lane) t32a[l] += __popc(mmm);
        __syncwarp();
    }
    // agent id per LOCAL agent, compacted in place (first[l] >= l, so a chunk never overwrites what a later one reads)
    for (int l0 = 0; l0 < n; l0 += 32) {
        int l = l0 + lane;
        int v = (l < n) ? c.agent[c.first[l]] : 0;
        __syncwarp();
        if (l < n) c.agent[l] = v;
    }
    __syncwarp();

    // ---- 3. per-agent occupancy masks + total timesteps; adjacency bit matrix by AND + ballot ----
    const int bw = (n + 31) >> 5;
    int ts_sum = 0;
    for (int l = lane; l < n; l += 32) {
        uint32_t mk[5] = { 0, 0, 0, 0, 0 };
        int t = 0;
        for (int q = c.aoff[l]; q < c.aoff[l + 1]; q++) {
            int p = c.apos[q]; int s = c.se[p] & 0xff, e = c.se[p] >> 8;
            int len = e - s + 1; t += len > 0 ? len : 0;
#pragma unroll
            for (int k = 0; k < 5; k++) mk[k] |= range_word(s, e, k);
        }
        c.ts[l] = (uint16_t)t; ts_sum += t;
        uint2 *dst = (uint2 *)(c.mask + 6 * l);
        dst[0] = make_uint2(mk[0], mk[1]); dst[1] = make_uint2(mk[2], mk[3]); dst[2] = make_uint2(mk[4], 0u);
    }
    ts_sum = __reduce_add_sync(FULLMASK, ts_sum);
    __syncwarp();
    int pc_sum = 0;
    for (int yb = 0; yb < bw; yb++) {
        const int y = yb * 32 + lane;
        uint2 m0 = make_uint2(0, 0), m1 = m0, m2 = m0;
        if (y < n) { const uint2 *src = (const uint2 *)(c.mask + 6 * y); m0 = src[0]; m1 = src[1]; m2 = src[2]; }
        for (int x0 = 0; x0 < n; x0 += 32) {
            const int xe = min(32, n - x0);
            uint32_t keep = 0;
            for (int xi = 0; xi < xe; xi++) {
                const uint2 *ox = (const uint2 *)(c.mask + 6 * (x0 + xi));
                uint2 a0 = ox[0], a1 = ox[1], a2 = ox[2];
                uint32_t hit = (m0.x & a0.x) | (m0.y & a0.y) | (m1.x & a1.x) | (m1.y & a1.y) | (m2.x & a2.x);
                uint32_t word = __ballot_sync(FULLMASK, hit != 0 && y != x0 + xi);
                if (lane == xi) keep = word;
            }
            // lane xi holds word yb of row x0 + xi: coalesced store into the word-major scratch
            if (lane < xe) {
                __stcg(&bits[yb * MM + x0 + lane], keep);
                const int pcv = __popc(keep);
                pc_sum += pcv;
                c.pc[x0 + lane] = (uint16_t)(yb == 0 ? pcv : (int)c.pc[x0 + lane] + pcv);
            }
        }
    }
    pc_sum = __reduce_add_sync(FULLMASK, pc_sum);
    __syncwarp();
    acc_pot += (unsigned)pc_sum / 2;
    if (pc_sum == 0) {                                    // contactnetwork.py:198
        if (lane == 0) g.slot_nrows[slot] = 0;
        __syncwarp();
        return;
    }

    // ---- 4. degrees (contactnetwork.py:210-269) ----
    const mtd_params *P = g.P;
    const long long max_degree = (long long)P->k_by_act[act] * n;
    const double avg_ts = (double)ts_sum / (double)n;
    const double avg_pc = (double)pc_sum / (double)n;     // every agent has an entry once n >= 2
    const double one_minus_mhd = 1 - g.day.mhd;
    const double lg = log((lane & 1) ? avg_pc : avg_ts);                  // both logarithms in one pass (even / odd lanes)
    const double log_avg_ts = __shfl_sync(FULLMASK, lg, 0), log_avg_pc = __shfl_sync(FULLMASK, lg, 1);
    for (int l = lane; l < n; l += 32) {
        double d = 0.0;
        int pcv = c.pc[l];
        if (pcv > 0) d = degree_f(g, P, c.agent[l], act, c.ts[l], pcv, avg_ts, avg_pc, log_avg_ts, log_avg_pc, one_minus_mhd);
        degf[l] = d;
    }
    __syncwarp();
    long long deg_sum = 0;
    if (lane == 0) deg_sum = (long long)rint(np_pairwise_sum(degf, n));
    deg_sum = __shfl_sync(FULLMASK, deg_sum, 0);
    // ---- 5. greedy order: sorted(zip(deg, id), reverse=True) (contactnetwork.py:283) ----
    // key[l] takes the slot of degf[l] (same lane reads the float degree, then writes the key)
    int Pn = 2;
    while (Pn < n) Pn <<= 1;
    int n_act = 0;                                         // agents with initial degree > 0 lead the order
    for (int l0 = 0; l0 < Pn; l0 += 32) {
        const int l = l0 + lane;
        int dg = 0;
        uint64_t key = ~0ull;
        if (l < n) {
            dg = degree_i(degf[l], deg_sum, max_degree, c.pc[l]);
            c.deg[l] = (int16_t)dg;
            key = ~(((uint64_t)(uint16_t)dg << 48) | ((uint64_t)(uint32_t)c.agent[l] << 16) | (uint64_t)l);
        }
        if (l < Pn) c.key[l] = key;
        const uint32_t al = __ballot_sync(FULLMASK, dg > 0);
        if (lane == 0) c.alive[l0 >> 5] = al;              // word of the alive mask (remaining degree > 0)
        n_act += __popc(al);
    }
    __syncwarp();
    warp_bitonic_sort(c.key, Pn, lane);
    uint16_t *ord = c.la;                                  // la is dead after phase 2
    for (int l = lane; l < n_act; l += 32) ord[l] = (uint16_t)((~c.key[l]) & 0xffff);
    __syncwarp();

    // ---- 6. greedy sampling + create_contacts (contactnetwork.py:287-301, 386-424) ----
    // Per main agent x: candidates = adjacency row of x AND alive (util.get_all_contacts_ids_by_id).  Philox mode: draw t
    // takes the r_t-th REMAINING candidate in canonical order, r_t = floor(u_t * (nc - t)).  That sequence is a Lehmer
    // code: lane t decodes its own pick's rank among the ORIGINAL candidates by walking the earlier draws backwards,
    // q += (q >= r_s) for s = t-1 .. 0, and selects that set bit -- all picks of the main agent in parallel.
    uint16_t *rarr = c.ts;                                 // draws r_t (ts is dead after phase 4)
    uint16_t *pickarr = c.pc;                              // replay: picked local agents (pc is dead after phase 4)
    uint32_t *stage = (uint32_t *)(sm + L::o_u);           // reserved rows (pairs of local agents); U is dead from here on
    constexpr int stage_cap = (int)(L::u_bytes / sizeof(uint32_t));
    RowRec *rows = g.scratch_rows + g.slot_rowoff[slot];
    const int row_cap = g.slot_rowbound[slot];
    const bool replay = g.day.mode == MTD_MODE_REPLAY;
    int n_rows = 0;
    uint32_t pre = (lane < bw && n_act > 0) ? __ldcg(&bits[lane * MM + ord[0]]) : 0u;
    for (int oi = 0; oi < n_act; oi++) {
        const int x = ord[oi];
        const uint32_t roww = pre;
        if (oi + 1 < n_act) pre = (lane < bw) ? __ldcg(&bits[lane * MM + ord[oi + 1]]) : 0u;
        const int dx = c.deg[x];
        if (dx <= 0) continue;
        const uint32_t cw = (lane < bw) ? (roww & c.alive[lane]) : 0u;
        const int cnt = __popc(cw);
        const int nc = __reduce_add_sync(FULLMASK, cnt);
        if (nc == 0) continue;                               // returned None (contactnetwork.py:294)
        const int d = dx < nc ? dx : nc;
        const int agx = c.agent[x];
        if (lane < bw) { c.candw[lane] = cw; c.ccnt[lane] = (uint32_t)cnt; }
        int made = d, spent = d;
        if (replay) {
            // the reference's selected partners (agent ids) -> local indices, validated against the candidate set
            long long tr = tape_find_sel(g.tape, cell, agx);
            if (tr < 0) { if (lane == 0) atomicAdd(&g.ctr->tape_miss, 1); continue; }
            const long long toff = g.tape.sel_off[tr];
            const int tc = g.tape.sel_off[tr + 1] - (int)toff;
            int npick = d;
            if (tc != d) { if (lane == 0) atomicAdd(&g.ctr->tape_miss, 1); if (tc < npick) npick = tc; }
            spent = npick;
            made = 0;
            __syncwarp();
            for (int t = 0; t < npick; t++) {
                const int pid = g.tape.sel_picks[toff + t];
                int y = -1;
                for (int l0 = 0; l0 < n; l0 += 32) {
                    const int l = l0 + lane;
                    const uint32_t hb = __ballot_sync(FULLMASK, l < n && c.agent[l] == pid);
                    if (hb) { y = l0 + __ffs(hb) - 1; break; }
                }
                const bool ok = y >= 0 && ((c.candw[y >> 5] >> (y & 31)) & 1u);
                __syncwarp();
                if (!ok) { if (lane == 0) atomicAdd(&g.ctr->tape_miss, 1); continue; }
                if (lane == 0) { c.candw[y >> 5] &= ~(1u << (y & 31)); pickarr[made] = (uint16_t)y; }
                made++;
                __syncwarp();
            }
        } else {
            for (int t = lane; t < d; t += 32) {
                uint32_t w4[4];
                philox4x32((uint32_t)g.day.day, (uint32_t)cell, (uint32_t)agx, (uint32_t)(t >> 2), g.day.seed, SITE_SELECT, w4);
                rarr[t] = (uint16_t)__umulhi(pick4(w4, t & 3), (uint32_t)(nc - t));
            }
        }
        __syncwarp();
        for (int t = lane; t < made; t += 32) {
            int y;
            if (replay) y = pickarr[t];
            else {
                int q = rarr[t];
                for (int s = t - 1; s >= 0; s--) q += (q >= (int)rarr[s]);
                int w = 0, cn = (int)c.ccnt[0];
                while (q >= cn) { q -= cn; w++; cn = (int)c.ccnt[w]; }
                y = w * 32 + nth_set_bit(c.candw[w], q);
            }
            // partner degree - 1; dead at 0 (delete_all_pairs_by_id).  Picks are distinct agents: no conflicts.
            const int nd = (int)c.deg[y] - 1;
            c.deg[y] = (int16_t)nd;
            if (nd == 0) atomicAnd(&c.alive[y >> 5], ~(1u << (y & 31)));
            // the row is only reserved here (local indices of the pair); its overlap scan runs after the greedy loop with
            // all 32 lanes on rows, not `made` lanes inside the sequential chain
            const int ri = n_rows + t;
            const uint32_t xy = (uint32_t)x | ((uint32_t)y << 16);
            if (ri >= row_cap) g.ctr->error = MTD_E_CAPACITY;
            else if (ri < stage_cap) stage[ri] = xy;
            else __stcg(&rows[ri].a1, (int32_t)xy);                // beyond the shared scratch (rare): through the row itself
        }
        if (lane == 0) {                                     // the main agent's degree is spent
            c.deg[x] = (int16_t)(d - spent);
            if (d - spent == 0) atomicAnd(&c.alive[x >> 5], ~(1u << (x & 31)));
        }
        n_rows += made;
        __syncwarp();
    }
    const int n_out = n_rows < row_cap ? n_rows : row_cap;
    if (lane == 0) g.slot_nrows[slot] = n_out;
    __syncwarp();
    // ---- 7. contact rows of the reserved pairs (contactnetwork.py:98-105, 160-196) ----
    for (int r = lane; r < n_out; r += 32) {
        const uint32_t xy = r < stage_cap ? stage[r] : (uint32_t)__ldcg(&rows[r].a1);
        const int x = (int)(xy & 0xffffu), y = (int)(xy >> 16);
        bool x_first; int st, en, dur;
        pair_row_lists(c, x, y, x_first, st, en, dur);
        RowRec rr;
        const int agx = c.agent[x], agy = c.agent[y];
        rr.a1 = x_first ? agx : agy; rr.a2 = x_first ? agy : agx;
        rr.sed = (uint32_t)st | ((uint32_t)en << 8) | ((uint32_t)dur << 16);
        rr.slot = slot;
        rows[r] = rr;
    }
    __syncwarp();
}

// Longest-processing-time-first order of one large class: a 300-interval cell keeps one warp busy for ~0.6 ms, so it
// must be among the first cells fetched, not the last (the tail decides the stage when a rank holds only a few
// thousand such cells).  One block: counting sort of the class list by interval count, descending, into the extra
// list region.  The order of cells inside a bin is arbitrary; every cell writes only its own rows.
__global__ void __launch_bounds__(1024) k1_lpt_kernel(DevCtx g, int cls, int out_region, int m_hi, int n_bins)
{
    __shared__ int hist[512];
    const int total = g.ctr->class_cells[cls];
    const int32_t *list = g.class_list + (int64_t)cls * g.class_stride;
    int32_t *out = g.class_list + (int64_t)out_region * g.class_stride;
    for (int b = threadIdx.x; b < n_bins; b += blockDim.x) hist[b] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        int b = m_hi - g.slot_m[list[i]];
        b = b < 0 ? 0 : (b >= n_bins ? n_bins - 1 : b);
        atomicAdd(&hist[b], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) { int acc = 0; for (int b = 0; b < n_bins; b++) { const int h = hist[b]; hist[b] = acc; acc += h; } }
    __syncthreads();
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int slot = list[i];
        int b = m_hi - g.slot_m[slot];
        b = b < 0 ? 0 : (b >= n_bins ? n_bins - 1 : b);
        out[atomicAdd(&hist[b], 1)] = slot;
    }
}

template <int MM, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, (MM == 128 ? 8 : 1)) k1_large_kernel(DevCtx g, int cls, int64_t bits_off, int list_cls)
{
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char *sm = smem + (size_t)warp * Lay<MM>::bytes;
    uint32_t *bits = g.k1_bits + bits_off + ((size_t)blockIdx.x * WARPS + warp) * Lay<MM>::bits_words;
    if (g.ctr->error) return;
    const int total = g.ctr->class_cells[cls];
    const int32_t *list = g.class_list + (int64_t)list_cls * g.class_stride;
    unsigned long long acc_pairs = 0, acc_pot = 0;
    while (true) {
        int w = 0;
        if (lane == 0) w = atomicAdd(&g.ctr->class_cursor[cls], 1);
        w = __shfl_sync(FULLMASK, w, 0);
        if (w >= total) break;
        process_cell_large<MM>(g, sm, bits, list[w], lane, acc_pairs, acc_pot);
    }
    if (lane == 0 && (acc_pairs | acc_pot)) {
        atomicAdd(&g.ctr->pairs_evaluated, acc_pairs);
        atomicAdd(&g.ctr->n_potential, acc_pot);
    }
}

// resident blocks per SM of one large-cell class (cap: the 1024 class is almost empty, keep its scratch small)
template <int MM, int WARPS>
int large_blocks_per_sm(int cap)
{
    static int per_sm = 0;
    if (per_sm == 0) {
        size_t smem = Lay<MM>::bytes * WARPS;
        cudaFuncSetAttribute(k1_large_kernel<MM, WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_large_kernel<MM, WARPS>, WARPS * 32, smem);
        if (per_sm < 1) per_sm = 1;
    }
    return per_sm > cap ? cap : per_sm;
}

template <int MM, int WARPS>
int launch_large(const DevCtx &c, int cls, cudaStream_t st, int sm_count, int cap, int64_t bits_off, int lpt_region = 0)
{
    int grid = sm_count * large_blocks_per_sm<MM, WARPS>(cap);
    const int64_t fit = (c.k1_bits_words - bits_off) / (int64_t)(Lay<MM>::bits_words * WARPS);
    if (grid > fit) grid = (int)fit;
    if (grid < 1) return 0;
    if (lpt_region) k1_lpt_kernel<<<1, 1024, 0, st>>>(c, cls, lpt_region, MM, 512);
    k1_large_kernel<MM, WARPS><<<grid, WARPS * 32, Lay<MM>::bytes * WARPS, st>>>(c, cls, bits_off, lpt_region ? lpt_region : cls);
    return lpt_region ? 2 : 1;
}


}  // namespace

// words of the adjacency scratch (DevCtx::k1_bits) the legacy launches need on a device with sm_count SMs
int64_t k1_legacy_bits_words(int sm_count)
{
    int64_t w5 = (int64_t)sm_count * large_blocks_per_sm<1024, 1>(2) * 1 * (int64_t)Lay<1024>::bits_words;
    int64_t w4 = (int64_t)sm_count * large_blocks_per_sm<320, 5>(64) * 5 * (int64_t)Lay<320>::bits_words;
    int64_t w3 = (int64_t)sm_count * large_blocks_per_sm<128, 4>(64) * 4 * (int64_t)Lay<128>::bits_words;
    return w5 + 2 * w4 + 2 * w3;       // one region per launch: the five launches may run side by side
}

// size classes 3..7 of k0_csr.cu (<= 64, 128, 192, 320, 1024 intervals) through the warp-per-cell kernels
int launch_k1_legacy(const DevCtx &c, cudaStream_t sa, cudaStream_t sb, int sm_count)
{
    const int64_t w5 = (int64_t)sm_count * large_blocks_per_sm<1024, 1>(2) * 1 * (int64_t)Lay<1024>::bits_words;
    const int64_t w4 = (int64_t)sm_count * large_blocks_per_sm<320, 5>(64) * 5 * (int64_t)Lay<320>::bits_words;
    const int64_t w3 = (int64_t)sm_count * large_blocks_per_sm<128, 4>(64) * 4 * (int64_t)Lay<128>::bits_words;
    int n = 0;
    n += launch_large<1024, 1>(c, 7, sa, sm_count, 2, 0);
    n += launch_large<320, 5>(c, 6, sa, sm_count, 64, w5);
    n += launch_large<320, 5>(c, 5, sb, sm_count, 64, w5 + w4);
    n += launch_large<128, 4>(c, 4, sb, sm_count, 64, w5 + 2 * w4);
    n += launch_large<128, 4>(c, 3, sb, sm_count, 64, w5 + 2 * w4 + w3);
    return n;
}
