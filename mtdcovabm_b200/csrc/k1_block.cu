// K1, cells of more than 32 intervals -- ContactNetwork.generate_contact_network (contactnetwork.py:134-315) and the row
// conversion of simulate_contact_network (contactnetwork.py:98-105), one thread BLOCK (2..16 warps) per cell.
//
// Everything of a cell lives in shared memory (cells of more than 1024 intervals: in a per-block global workspace, same
// code).  The potential-contact graph is a bit matrix over interval POSITIONS (list order of the cell):
//   * the cell's CSR segment is staged by one bulk-async copy (cp.async.bulk + mbarrier), sorted back into emission
//     order, and the agents are numbered by first appearance (contactnetwork.py:161-163);
//   * S[t] = {j : start_j <= t} and E[t] = {j : end_j >= t} are bit sets over positions for every timestep t, built by one
//     atomicOr per interval and a prefix / suffix OR over t.  Interval i overlaps exactly the intervals S[end_i] & E[start_i]
//     (inclusive bounds, contactnetwork.py:426-432): the row of an agent is the OR over its intervals -- O(m^2 / 32) word
//     operations with two loads each instead of an O(m^2) pair loop;
//   * bits of an agent's later intervals are folded onto its first position, so a row ends up as a bit set over FIRST
//     positions = agents in order of first appearance, the canonical candidate order of the Philox selection;
//   * degrees in fp64 (numpy's pairwise summation order, contactnetwork.py:210-269), greedy order by bitonic sort;
//   * the sequential greedy sampling (contactnetwork.py:283-301, 386-424) is run by warp 0 on bit rows with an alive mask;
//     the first Philox block of every agent is drawn beforehand by all warps, so the chain only scales it;
//   * the contact rows of the selected pairs are built afterwards by all threads.
#include "common.cuh"
#include <stdlib.h>

namespace {

#define K1_CELLS_PER_BLOCK 2    // cells a block takes before it gives its SM slot back (while blocks of its grid still wait)
#define K1_GRID_ROUNDS 8        // grid = resident blocks x this
#define K1_TS_ROWS 160          // timestep domain of the S/E tables (the small-cell kernel's 5 x 32-bit masks)

__host__ __device__ constexpr int pow2_ge(int x) { int p = 1; while (p < x) p <<= 1; return p; }
__host__ __device__ constexpr size_t al16(size_t x) { return (x + 15) & ~(size_t)15; }

// byte offsets of one cell's working set for a capacity of mm intervals (mm a multiple of 32)
struct Lay {
    int mm, bw;
    size_t o_agent, o_se, o_la, o_apos, o_aoff, o_first, o_pc, o_deg, o_ts, o_ord, o_F, o_alive, o_candw, o_cincl, o_tmpw,
           o_u, o_bits, bytes;
    __host__ __device__ constexpr Lay(int mm_) : mm(mm_), bw(mm_ / 32), o_agent(0), o_se(0), o_la(0), o_apos(0), o_aoff(0),
        o_first(0), o_pc(0), o_deg(0), o_ts(0), o_ord(0), o_F(0), o_alive(0), o_candw(0), o_cincl(0), o_tmpw(0), o_u(0),
        o_bits(0), bytes(0)
    {
        size_t o = 0;
        o_u = o;     o += al16((size_t)40 * mm);          // phase-aliased scratch, see process_cell
        o_agent = o; o += al16((size_t)4 * mm);           // i32 agent id per position, then per local agent
        o_se = o;    o += al16((size_t)2 * mm);           // u16 start | end << 8 per position
        o_la = o;    o += al16((size_t)2 * mm);           // u16 local agent of a position
        o_apos = o;  o += al16((size_t)2 * mm);           // u16 positions grouped by agent, ascending
        o_aoff = o;  o += al16((size_t)2 * (mm + 2));     // u16 offsets into apos
        o_first = o; o += al16((size_t)2 * mm);           // u16 first position of a local agent
        o_pc = o;    o += al16((size_t)2 * mm);           // u16 potential contacts; replay picks during the chain
        o_deg = o;   o += al16((size_t)2 * mm);           // i16 remaining degree
        o_ts = o;    o += al16((size_t)2 * mm);           // u16 total timesteps; the draws r_t during the chain
        o_ord = o;   o += al16((size_t)2 * mm);           // u16 greedy order
        o_F = o;     o += al16((size_t)4 * (bw + 1));     // first-position bit set
        o_alive = o; o += al16((size_t)4 * (bw + 1));     // first positions of agents with remaining degree > 0
        o_candw = o; o += al16((size_t)4 * (bw + 1));     // candidates of the current main agent
        o_cincl = o; o += al16((size_t)4 * (bw + 1));     // inclusive popcount prefix of candw
        o_tmpw = o;  o += al16((size_t)4 * (bw + 2));     // chunk totals of the block scans
        o_bits = o;  o += al16((size_t)4 * mm * bw);      // adjacency rows [local agent][word]
        bytes = o;
    }
};

struct Cell {
    unsigned char *u;
    int32_t *agent; uint16_t *se, *la, *apos, *aoff, *first, *pc; int16_t *deg; uint16_t *ts, *ord;
    uint32_t *F, *alive, *candw, *cincl, *tmpw, *bits;
};

__device__ __forceinline__ Cell make_cell(unsigned char *sm, const Lay &L)
{
    Cell c;
    c.u = sm + L.o_u; c.agent = (int32_t *)(sm + L.o_agent); c.se = (uint16_t *)(sm + L.o_se); c.la = (uint16_t *)(sm + L.o_la);
    c.apos = (uint16_t *)(sm + L.o_apos); c.aoff = (uint16_t *)(sm + L.o_aoff); c.first = (uint16_t *)(sm + L.o_first);
    c.pc = (uint16_t *)(sm + L.o_pc); c.deg = (int16_t *)(sm + L.o_deg); c.ts = (uint16_t *)(sm + L.o_ts);
    c.ord = (uint16_t *)(sm + L.o_ord); c.F = (uint32_t *)(sm + L.o_F); c.alive = (uint32_t *)(sm + L.o_alive);
    c.candw = (uint32_t *)(sm + L.o_candw); c.cincl = (uint32_t *)(sm + L.o_cincl); c.tmpw = (uint32_t *)(sm + L.o_tmpw);
    c.bits = (uint32_t *)(sm + L.o_bits);
    return c;
}

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

// contactnetwork.py:225-252: float degree of one agent (one 16-byte gather: {soc_rate, age_bracket})
__device__ __forceinline__ double degree_f(const DevCtx &g, const mtd_params *P, int ag, int act, int ts, int pc, double avg_ts,
                                           double avg_pc, double log_avg_ts, double log_avg_pc, double one_minus_mhd)
{
    const int4 v = __ldg((const int4 *)(g.k1_attr + ag));
    const double soc = __hiloint2double(v.y, v.x);
    int ab = v.z;
    if (ab < 0) ab += 16;                                   // numpy negative index
    const double cmv = P->cm[ab][act];
    double tsm = 1.0, pcm = 1.0;
    if ((double)ts != avg_ts && avg_ts > 1) tsm = log_int(g, ts) / log_avg_ts;       // math.log(x, base)
    if ((double)pc != avg_pc && avg_pc > 1) pcm = log_int(g, pc) / log_avg_pc;
    return cmv * tsm * pcm * soc * one_minus_mhd;
}

__device__ __forceinline__ int degree_i(double d, long long deg_sum, long long max_degree, int pc)
{   // contactnetwork.py:254-269 (Python round = rint)
    long long r;
    if (deg_sum > max_degree) r = (long long)rint((d / (double)deg_sum) * (double)max_degree);
    else r = (long long)rint(d);
    return (int)(r > pc ? pc : r);
}

template <int CW> __device__ __forceinline__ void bsync() { if (CW == 1) __syncwarp(); else __syncthreads(); }

template <int CW>
__device__ __forceinline__ int block_sum(int v, int *s_red, int lane, int wid)
{
    v = __reduce_add_sync(FULLMASK, v);
    if (CW == 1) return v;
    if (lane == 0) s_red[wid] = v;
    __syncthreads();
    int t = 0;
#pragma unroll
    for (int w = 0; w < CW; w++) t += s_red[w];
    __syncthreads();
    return t;
}

__device__ __forceinline__ int warp_incl_scan(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULLMASK, v, o); if (lane >= o) v += t; }
    return v;
}

template <int CW>
__device__ void block_bitonic_sort(uint64_t *k, int P2, int tid)
{
    for (int size = 2; size <= P2; size <<= 1) {
        for (int j = size >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < (P2 >> 1); t += CW * 32) {
                const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const int p = i | j;
                const bool asc = ((i & size) == 0);
                const uint64_t a = k[i], b = k[p];
                if ((a > b) == asc) { k[i] = b; k[p] = a; }
            }
            bsync<CW>();
        }
    }
}

// numpy float64 add.reduce (pairwise summation: 8 accumulators per block of <= 128, halves above), executed by one WARP:
// the eight accumulator chains of a block run on lanes 0..7.  Every lane returns the sum.
__device__ double np_pairwise_sum_warp(const double *a, int n, int lane)
{
    if (n < 8) {
        double r = -0.0;
        for (int i = 0; i < n; i++) r += a[i];
        return r;
    }
    if (n <= 128) {
        const int nfull = n - (n % 8);
        double r = (lane < 8) ? a[lane] : 0.0;
        if (lane < 8) for (int i = 8 + lane; i < nfull; i += 8) r += a[i];
        const double r0 = __shfl_sync(FULLMASK, r, 0), r1 = __shfl_sync(FULLMASK, r, 1), r2 = __shfl_sync(FULLMASK, r, 2),
                     r3 = __shfl_sync(FULLMASK, r, 3), r4 = __shfl_sync(FULLMASK, r, 4), r5 = __shfl_sync(FULLMASK, r, 5),
                     r6 = __shfl_sync(FULLMASK, r, 6), r7 = __shfl_sync(FULLMASK, r, 7);
        double res = ((r0 + r1) + (r2 + r3)) + ((r4 + r5) + (r6 + r7));
        for (int i = nfull; i < n; i++) res += a[i];
        return res;
    }
    int n2 = n / 2;
    n2 -= n2 % 8;
    const double left = np_pairwise_sum_warp(a, n2, lane);
    return left + np_pairwise_sum_warp(a + n2, n - n2, lane);
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
    int b_min = 0x7fffffff, b_max = -1, b_start = 0, b_end = 0;
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
                const int kb = (pb << 15) | pa;                     // positions < 32768
                if (kb < b_min) { b_min = kb; b_start = lo; }
                if (kb > b_max) { b_max = kb; b_end = hi; }
            }
        }
    }
    x_first = fx < fy;
    start = x_first ? a_start : b_start;
    end = x_first ? a_end : b_end;
}

// ---- bulk-async (TMA) staging of the CSR segment
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bulk_load(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar)
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");        // earlier generic-proxy accesses of the buffer first
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void l2_prefetch(const void *src_gmem, uint32_t bytes)
{
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(src_gmem), "r"(bytes) : "memory");
}

struct BlockShared {
    uint64_t mbar;
    double lg[2];
    long long deg_sum;
    int red[16];
    int w, n_act, sorted;
};

// GWS: the working set lives in a global workspace (cells above the shared-memory classes); plain loads instead of TMA
template <int CW, bool GWS>
__device__ void process_cell(const DevCtx &g, const Lay &L, unsigned char *sm, BlockShared &bs, uint32_t &mbar_phase, int slot,
                             int next_slot, unsigned long long &acc_pairs, unsigned long long &acc_pot, unsigned long long &acc_inc,
                             bool skip_chain, bool split)
{
    constexpr int T = CW * 32;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const Cell c = make_cell(sm, L);
    const int m = g.slot_m[slot];
    const int64_t base = g.slot_off[slot];
    const int cell = g.slot_cell[slot];
    const int act = g.slot_act[slot];
    const int bwm = (m + 31) >> 5;                          // words per bit row of THIS cell
    int P2 = 2;
    while (P2 < m) P2 <<= 1;

    // scratch U, phase by phase:
    //   1   staging IvRec[m] at 0 | sort keys u64[P2] at 16 mm
    //   2   hash u32[2 P2] at 0   | t32a u32[P2], t32b u32[P2] at 16 mm
    //   3   S table u32[160][bwm] at 0 | E table at 20 mm
    //   4-6 degf f64[n] / order keys u64[Pn] at 0 | first Philox block uint4[n] at 16 mm | reserved rows u32[2 mm] at 32 mm
    IvRec *staging = (IvRec *)c.u;
    uint64_t *key = (uint64_t *)(c.u + (size_t)16 * L.mm);
    uint32_t *hash = (uint32_t *)c.u;
    uint32_t *t32a = (uint32_t *)key, *t32b = t32a + P2;
    uint32_t *tabS = (uint32_t *)c.u, *tabE = (uint32_t *)(c.u + (size_t)20 * L.mm);
    double *degf = (double *)c.u;
    uint4 *draws4 = (uint4 *)(c.u + (size_t)16 * L.mm);
    uint32_t *stage = (uint32_t *)(c.u + (size_t)32 * L.mm);
    const int stage_cap = 2 * L.mm;

    // ---- 1. CSR segment -> emission (list) order ----
    const IvRec *seg = g.csr + base;
    if (!GWS) {
        if (tid == 0) {
            bulk_load(staging, seg, (uint32_t)m * (uint32_t)sizeof(IvRec), &bs.mbar);
            if (next_slot >= 0) l2_prefetch(g.csr + g.slot_off[next_slot], (uint32_t)g.slot_m[next_slot] * (uint32_t)sizeof(IvRec));
        }
        mbar_wait(&bs.mbar, mbar_phase);
        mbar_phase ^= 1u;
        seg = staging;
    }
    bool sorted = true;
    for (int p = tid; p < P2; p += T) {
        uint64_t k = ~0ull;
        if (p < m) {
            const uint32_t o = seg[p].order;
            k = ((uint64_t)o << 32) | (uint32_t)p;
            if (p > 0 && seg[p - 1].order > o) sorted = false;
        }
        key[p] = k;
    }
    if (CW == 1) sorted = __all_sync(FULLMASK, sorted);
    else sorted = __syncthreads_and(sorted);
    if (!sorted) block_bitonic_sort<CW>(key, P2, tid);
    for (int p = tid; p < m; p += T) {
        const IvRec r = seg[(uint32_t)key[p]];
        c.agent[p] = r.agent; c.se[p] = (uint16_t)(r.se & 0xffffu);
    }
    bsync<CW>();

    // ---- 2. local agent index = order of first appearance (contactnetwork.py:161-163); per-agent interval lists ----
    const int H = 2 * P2;
    for (int h = tid; h < H; h += T) hash[h] = 0;
    for (int p = tid; p < P2; p += T) { t32a[p] = 0xffffffffu; t32b[p] = 0; }
    bsync<CW>();
    for (int p = tid; p < m; p += T) {                   // representative position of every agent
        const int ag = c.agent[p];
        uint32_t h = ((uint32_t)ag * 2654435761u) & (H - 1);
        int rep;
        while (true) {
            const uint32_t v = atomicCAS(&hash[h], 0u, (uint32_t)p + 1u);
            if (v == 0) { rep = p; break; }
            if (c.agent[v - 1] == ag) { rep = (int)v - 1; break; }
            h = (h + 1) & (H - 1);
        }
        c.la[p] = (uint16_t)rep;
    }
    bsync<CW>();
    for (int p = tid; p < m; p += T) atomicMin(&t32a[c.la[p]], (uint32_t)p);      // first position per representative
    bsync<CW>();
    for (int ch = wid; ch < bwm; ch += CW) {             // F = bit set of the first positions
        const int p = ch * 32 + lane;
        const bool isfirst = (p < m) && (t32a[c.la[p]] == (uint32_t)p);
        const uint32_t bm = __ballot_sync(FULLMASK, isfirst);
        if (lane == 0) { c.F[ch] = bm; c.tmpw[ch] = (uint32_t)__popc(bm); }
    }
    bsync<CW>();
    if (wid == 0) {                                       // exclusive prefix of the chunk counts
        int run = 0;
        for (int c0 = 0; c0 < bwm; c0 += 32) {
            const int ch = c0 + lane;
            const int v = ch < bwm ? (int)c.tmpw[ch] : 0;
            const int s = warp_incl_scan(v, lane);
            if (ch < bwm) c.tmpw[ch] = (uint32_t)(run + s - v);
            run += __shfl_sync(FULLMASK, s, 31);
        }
        if (lane == 0) c.tmpw[bwm] = (uint32_t)run;
    }
    bsync<CW>();
    const int n = (int)c.tmpw[bwm];
    for (int p = tid; p < m; p += T) {                   // number the first positions in position order
        const uint32_t fw = c.F[p >> 5];
        if ((fw >> (p & 31)) & 1u) {
            const int l = (int)c.tmpw[p >> 5] + __popc(fw & ((1u << (p & 31)) - 1u));
            c.first[l] = (uint16_t)p; c.pc[p] = (uint16_t)l;     // pc temporarily maps first position -> local index
        }
    }
    bsync<CW>();
    for (int p = tid; p < m; p += T) {
        const int l = c.pc[t32a[c.la[p]]];
        c.la[p] = (uint16_t)l;
        atomicAdd(&t32b[l], 1u);
    }
    bsync<CW>();
    int sq = 0;
    for (int ch = wid; ch < ((n + 31) >> 5); ch += CW) {  // exclusive scan of the per-agent interval counts -> aoff
        const int l = ch * 32 + lane;
        const int v = l < n ? (int)t32b[l] : 0;
        sq += v * v;
        const int s = warp_incl_scan(v, lane);
        if (l < n) c.aoff[l] = (uint16_t)(s - v);
        if (lane == 31) c.tmpw[ch] = (uint32_t)s;
    }
    sq = block_sum<CW>(sq, bs.red, lane, wid);
    bsync<CW>();
    if (wid == 0) {
        const int nch = (n + 31) >> 5;
        int run = 0;
        for (int c0 = 0; c0 < nch; c0 += 32) {
            const int ch = c0 + lane;
            const int v = ch < nch ? (int)c.tmpw[ch] : 0;
            const int s = warp_incl_scan(v, lane);
            if (ch < nch) c.tmpw[ch] = (uint32_t)(run + s - v);
            run += __shfl_sync(FULLMASK, s, 31);
        }
    }
    bsync<CW>();
    for (int l = tid; l < n; l += T) {
        const int o = (int)c.aoff[l] + (int)c.tmpw[l >> 5];
        c.aoff[l] = (uint16_t)o; t32a[l] = (uint32_t)o;  // t32a: fill cursor
    }
    if (tid == 0) { c.aoff[n] = (uint16_t)m; acc_pairs += (unsigned long long)((long long)m * m - sq) / 2; acc_inc += (unsigned)n; }
    bsync<CW>();
    for (int p = tid; p < m; p += T) c.apos[atomicAdd(&t32a[c.la[p]], 1u)] = (uint16_t)p;
    bsync<CW>();
    for (int l = tid; l < n; l += T) {                   // positions of an agent in ascending order (short lists)
        const int a0 = c.aoff[l], a1 = c.aoff[l + 1];
        for (int i = a0 + 1; i < a1; i++) {
            const uint16_t v = c.apos[i];
            int j = i - 1;
            while (j >= a0 && c.apos[j] > v) { c.apos[j + 1] = c.apos[j]; j--; }
            c.apos[j + 1] = v;
        }
        t32b[l] = (uint32_t)c.agent[c.first[l]];         // agent id per LOCAL agent, through a temporary
    }
    bsync<CW>();
    for (int l = tid; l < n; l += T) c.agent[l] = (int32_t)t32b[l];
    bsync<CW>();

    // ---- 3. adjacency rows over positions from the start / end tables ----
    for (int i = tid; i < 2 * K1_TS_ROWS * bwm; i += T) {
        if (i < K1_TS_ROWS * bwm) tabS[i] = 0u; else tabE[i - K1_TS_ROWS * bwm] = 0u;
    }
    bsync<CW>();
    for (int p = tid; p < m; p += T) {
        const int s = c.se[p] & 0xff;
        int e = c.se[p] >> 8;
        if (s <= e && s < K1_TS_ROWS) {                   // an empty or out-of-domain interval overlaps nothing
            if (e > K1_TS_ROWS - 1) e = K1_TS_ROWS - 1;
            atomicOr(&tabS[s * bwm + (p >> 5)], 1u << (p & 31));
            atomicOr(&tabE[e * bwm + (p >> 5)], 1u << (p & 31));
        }
    }
    bsync<CW>();
    {   // S[t] |= S[t-1] (prefix over t), E[t] |= E[t+1] (suffix over t): local scans over segments of 16 timesteps, then
        // every segment ORs in the boundary words of the segments before (after) it.  A boundary word may already hold
        // its own carry when it is read -- harmless: that carry is a subset of the boundary words that are ORed in anyway.
        constexpr int SEG = 16, NSEG = K1_TS_ROWS / SEG;
        const int tasks = 2 * bwm * NSEG;
        for (int i = tid; i < tasks; i += T) {
            const int col = i % bwm, r = i / bwm, sgi = r % NSEG, which = r / NSEG;
            uint32_t acc = 0;
            if (which == 0) for (int t = sgi * SEG; t < sgi * SEG + SEG; t++) { acc |= tabS[t * bwm + col]; tabS[t * bwm + col] = acc; }
            else for (int t = sgi * SEG + SEG - 1; t >= sgi * SEG; t--) { acc |= tabE[t * bwm + col]; tabE[t * bwm + col] = acc; }
        }
        bsync<CW>();
        for (int i = tid; i < tasks; i += T) {
            const int col = i % bwm, r = i / bwm, sgi = r % NSEG, which = r / NSEG;
            uint32_t carry = 0;
            if (which == 0) for (int s2 = 0; s2 < sgi; s2++) carry |= tabS[(s2 * SEG + SEG - 1) * bwm + col];
            else for (int s2 = sgi + 1; s2 < NSEG; s2++) carry |= tabE[(s2 * SEG) * bwm + col];
            uint32_t *tab = which == 0 ? tabS : tabE;
            if (carry) for (int t = sgi * SEG; t < sgi * SEG + SEG; t++) tab[t * bwm + col] |= carry;
        }
    }
    bsync<CW>();
    int ts_sum = 0;
    {   // row of agent l, word w = OR over its intervals i of S[end_i][w] & E[start_i][w]
        const uint32_t total = (uint32_t)n * (uint32_t)bwm;
        const uint32_t rcp = 0xffffffffu / (uint32_t)bwm + 1u;
        for (uint32_t i = tid; i < total; i += T) {
            uint32_t l = GWS ? i / (uint32_t)bwm : __umulhi(i, rcp);
            uint32_t w = i - l * (uint32_t)bwm;
            if (!GWS && w >= (uint32_t)bwm) { l -= 1; w += (uint32_t)bwm; }
            uint32_t acc = 0;
            int tsl = 0;
            for (int q = c.aoff[l]; q < c.aoff[l + 1]; q++) {
                const int p = c.apos[q];
                const int s = c.se[p] & 0xff;
                int e = c.se[p] >> 8;
                const int len = e - s + 1;
                tsl += len > 0 ? len : 0;
                if (s <= e && s < K1_TS_ROWS) {
                    if (e > K1_TS_ROWS - 1) e = K1_TS_ROWS - 1;
                    acc |= tabS[e * bwm + w] & tabE[s * bwm + w];
                }
            }
            c.bits[l * bwm + w] = acc;
            if (w == 0) { c.ts[l] = (uint16_t)tsl; ts_sum += tsl; }
        }
    }
    bsync<CW>();
    {   // fold the bits of later intervals onto their agent's first position
        const uint32_t total = (uint32_t)n * (uint32_t)bwm;
        const uint32_t rcp = 0xffffffffu / (uint32_t)bwm + 1u;
        for (uint32_t i = tid; i < total; i += T) {
            uint32_t l = GWS ? i / (uint32_t)bwm : __umulhi(i, rcp);
            uint32_t w = i - l * (uint32_t)bwm;
            if (!GWS && w >= (uint32_t)bwm) { l -= 1; w += (uint32_t)bwm; }
            for (uint32_t nf = c.bits[l * bwm + w] & ~c.F[w]; nf; nf &= nf - 1) {
                const int p = (int)w * 32 + __ffs(nf) - 1;
                const int q = c.first[c.la[p]];
                atomicOr(&c.bits[l * bwm + (q >> 5)], 1u << (q & 31));
            }
        }
    }
    bsync<CW>();
    int pc_sum = 0;
    for (int l = tid; l < n; l += T) {                   // keep first positions only, drop the agent itself; count
        const int f = c.first[l];
        int cnt = 0;
        for (int w = 0; w < bwm; w++) {
            uint32_t v = c.bits[l * bwm + w] & c.F[w];
            if ((f >> 5) == w) v &= ~(1u << (f & 31));
            c.bits[l * bwm + w] = v;
            cnt += __popc(v);
        }
        c.pc[l] = (uint16_t)cnt;
        pc_sum += cnt;
    }
    ts_sum = block_sum<CW>(ts_sum, bs.red, lane, wid);
    pc_sum = block_sum<CW>(pc_sum, bs.red, lane, wid);
    if (tid == 0) acc_pot += (unsigned)pc_sum / 2;
    if (pc_sum == 0) {                                    // contactnetwork.py:198
        if (tid == 0) {
            g.slot_nrows[slot] = 0;
            if (split) *(int4 *)(g.k1_rec + (size_t)MTD_K1_REC_STRIDE * (size_t)base) = make_int4(0, n, m, bwm);    // nothing for the chain kernel to do
        }
        bsync<CW>();
        return;
    }

    // ---- 4. degrees (contactnetwork.py:210-269) ----
    const mtd_params *P = g.P;
    const long long max_degree = (long long)P->k_by_act[act] * n;
    const double avg_ts = (double)ts_sum / (double)n;
    const double avg_pc = (double)pc_sum / (double)n;     // every agent has an entry once n >= 2
    const double one_minus_mhd = 1 - g.day.mhd;
    if (tid < 2) bs.lg[tid] = log(tid ? avg_pc : avg_ts);
    bsync<CW>();
    const double log_avg_ts = bs.lg[0], log_avg_pc = bs.lg[1];
    const bool replay = g.day.mode == MTD_MODE_REPLAY;
    for (int l = tid; l < n; l += T) {
        double d = 0.0;
        const int pcv = c.pc[l];
        if (pcv > 0) {
            d = degree_f(g, P, c.agent[l], act, c.ts[l], pcv, avg_ts, avg_pc, log_avg_ts, log_avg_pc, one_minus_mhd);
            if (!replay && !split) {                      // the first four draws of the agent, should it become a main agent
                uint32_t w4[4];
                philox4x32((uint32_t)g.day.day, (uint32_t)cell, (uint32_t)c.agent[l], 0u, g.day.seed, SITE_SELECT, w4);
                draws4[l] = make_uint4(w4[0], w4[1], w4[2], w4[3]);
            }
        }
        degf[l] = d;
    }
    for (int w = tid; w < bwm; w += T) c.alive[w] = 0u;
    bsync<CW>();
    if (wid == 0) {
        const double s = np_pairwise_sum_warp(degf, n, lane);
        if (lane == 0) bs.deg_sum = (long long)rint(s);
    }
    bsync<CW>();
    const long long deg_sum = bs.deg_sum;

    // ---- 5. greedy order: sorted(zip(deg, id), reverse=True) (contactnetwork.py:283) ----
    uint64_t *okey = (uint64_t *)c.u;                      // okey[l] takes the slot of degf[l]: read, then written by the same thread
    int Pn = 2;
    while (Pn < n) Pn <<= 1;
    int n_act = 0;                                         // agents with initial degree > 0 lead the order
    for (int l = tid; l < Pn; l += T) {
        uint64_t k = ~0ull;
        if (l < n) {
            const int dg = degree_i(degf[l], deg_sum, max_degree, c.pc[l]);
            c.deg[l] = (int16_t)dg;
            k = ~(((uint64_t)(uint16_t)dg << 48) | ((uint64_t)(uint32_t)c.agent[l] << 16) | (uint64_t)l);
            if (dg > 0) { n_act++; const int f = c.first[l]; atomicOr(&c.alive[f >> 5], 1u << (f & 31)); }
        }
        okey[l] = k;
    }
    n_act = block_sum<CW>(n_act, bs.red, lane, wid);
    block_bitonic_sort<CW>(okey, Pn, tid);
    for (int l = tid; l < n_act; l += T) c.ord[l] = (uint16_t)((~okey[l]) & 0xffff);
    bsync<CW>();

    if (split) {
        // ---- 5b. split path: leave the cell's record in HBM for the thread-per-cell chain kernel (k1_chain.cu) and for
        //          K2's row builder (K1Rec, common.cuh); the sequential sampling does not run in this kernel
        unsigned char *rec = g.k1_rec + (size_t)MTD_K1_REC_STRIDE * (size_t)base;
        const K1Rec R(m, n, bwm);
        uint32_t *mains = (uint32_t *)(rec + R.o_main);
        const uint32_t mw = R.main_words;
        if (tid == 0) *(int4 *)rec = make_int4(skip_chain ? 0 : n_act, n, m, bwm);
        for (int w = tid; w < (int)R.bw2; w += T) ((uint32_t *)(rec + R.o_alive))[w] = w < bwm ? c.alive[w] : 0u;
        for (uint32_t i = tid; i < (uint32_t)n_act * mw; i += T) {
            const uint32_t oi = i / mw, k = i - oi * mw;
            const int x = c.ord[oi];
            uint32_t v;
            if (k == 0) v = (uint32_t)x | ((uint32_t)c.first[x] << 16);
            else if (k == 1) v = (uint32_t)c.agent[x];
            else if (k < 6) continue;                      // the Philox block is written below
            else v = (int)(k - 6) < bwm ? c.bits[x * bwm + (k - 6)] : 0u;
            mains[i] = v;
        }
        if (!replay)                                       // the first four draws of every main agent (one Philox block)
            for (int oi = tid; oi < n_act; oi += T) {
                uint32_t w4[4];
                philox4x32((uint32_t)g.day.day, (uint32_t)cell, (uint32_t)c.agent[c.ord[oi]], 0u, g.day.seed, SITE_SELECT, w4);
                uint32_t *d4 = mains + (size_t)oi * mw + 2;
                d4[0] = w4[0]; d4[1] = w4[1]; d4[2] = w4[2]; d4[3] = w4[3];
            }
        int32_t *r_agent = (int32_t *)(rec + R.o_agent);
        uint16_t *r_deg = (uint16_t *)(rec + R.o_deg), *r_se = (uint16_t *)(rec + R.o_se), *r_la = (uint16_t *)(rec + R.o_la),
                 *r_apos = (uint16_t *)(rec + R.o_apos), *r_aoff = (uint16_t *)(rec + R.o_aoff);
        for (int l = tid; l < n; l += T) { r_agent[l] = c.agent[l]; r_aoff[l] = c.aoff[l]; }
        if (tid == 0) r_aoff[n] = c.aoff[n];
        for (int p = tid; p < m; p += T) {
            const int l = c.la[p];
            const int dg = c.deg[l];
            r_deg[p] = (uint16_t)((c.first[l] == p && dg > 0) ? dg : 0);
            r_se[p] = c.se[p]; r_la[p] = (uint16_t)l; r_apos[p] = c.apos[p];
        }
        if (tid == 0) g.slot_nrows[slot] = 0;            // the chain kernel sets the count
        bsync<CW>();
        return;
    }

    // ---- 6. greedy sampling + create_contacts (contactnetwork.py:287-301, 386-424), warp 0 ----
    // Per main agent x: candidates = adjacency row of x AND alive (util.get_all_contacts_ids_by_id).  Philox mode: draw t
    // takes the r_t-th REMAINING candidate in canonical order, r_t = floor(u_t * (nc - t)).  That sequence is a Lehmer
    // code: lane t decodes its own pick's rank among the ORIGINAL candidates by walking the earlier draws backwards,
    // q += (q >= r_s) for s = t-1 .. 0, and selects that set bit -- all picks of the main agent in parallel.
    RowRec *rows = g.scratch_rows + g.slot_rowoff[slot];
    const int row_cap = g.slot_rowbound[slot];
    int n_rows = 0;
    if (skip_chain) n_act = 0;                              // timing experiment (MTD_K1_SKIPCHAIN=1): phases 1-5 only, no rows
    if (wid == 0) {
        uint16_t *rarr = c.ts;                             // draws r_t (ts is dead after phase 4)
        uint16_t *pickarr = c.pc;                          // replay: picked local agents (pc is dead after phase 5)
        for (int oi = 0; oi < n_act; oi++) {
            const int x = c.ord[oi];
            const int dx = c.deg[x];
            if (dx <= 0) continue;
            int nc = 0;
            for (int w0 = 0; w0 < bwm; w0 += 32) {         // candidate words + inclusive popcount prefix
                const int w = w0 + lane;
                const uint32_t cw = w < bwm ? (c.bits[x * bwm + w] & c.alive[w]) : 0u;
                const int s = warp_incl_scan(__popc(cw), lane);
                if (w < bwm) { c.candw[w] = cw; c.cincl[w] = (uint32_t)(nc + s); }
                nc += __shfl_sync(FULLMASK, s, 31);
            }
            if (nc == 0) continue;                           // returned None (contactnetwork.py:294)
            const int d = dx < nc ? dx : nc;
            const int agx = c.agent[x];
            int made = d, spent = d;
            __syncwarp();
            if (replay) {
                // the reference's selected partners (agent ids) -> local indices, validated against the candidate set
                const long long tr = tape_find_sel(g.tape, cell, agx);
                if (tr < 0) { if (lane == 0) atomicAdd(&g.ctr->tape_miss, 1); continue; }
                const long long toff = g.tape.sel_off[tr];
                const int tc = g.tape.sel_off[tr + 1] - (int)toff;
                int npick = d;
                if (tc != d) { if (lane == 0) atomicAdd(&g.ctr->tape_miss, 1); if (tc < npick) npick = tc; }
                spent = npick;
                made = 0;
                for (int t = 0; t < npick; t++) {
                    const int pid = g.tape.sel_picks[toff + t];
                    int y = -1;
                    for (int l0 = 0; l0 < n; l0 += 32) {
                        const int l = l0 + lane;
                        const uint32_t hb = __ballot_sync(FULLMASK, l < n && c.agent[l] == pid);
                        if (hb) { y = l0 + __ffs(hb) - 1; break; }
                    }
                    const int f = y >= 0 ? (int)c.first[y] : 0;
                    const bool ok = y >= 0 && ((c.candw[f >> 5] >> (f & 31)) & 1u);
                    __syncwarp();
                    if (!ok) { if (lane == 0) atomicAdd(&g.ctr->tape_miss, 1); continue; }
                    if (lane == 0) { c.candw[f >> 5] &= ~(1u << (f & 31)); pickarr[made] = (uint16_t)y; }
                    made++;
                    __syncwarp();
                }
            } else {
                const uint4 dr = draws4[x];
                for (int t = lane; t < d; t += 32) {
                    uint32_t word;
                    if (t < 4) word = t == 0 ? dr.x : (t == 1 ? dr.y : (t == 2 ? dr.z : dr.w));
                    else {
                        uint32_t w4[4];
                        philox4x32((uint32_t)g.day.day, (uint32_t)cell, (uint32_t)agx, (uint32_t)(t >> 2), g.day.seed, SITE_SELECT, w4);
                        word = pick4(w4, t & 3);
                    }
                    rarr[t] = (uint16_t)__umulhi(word, (uint32_t)(nc - t));
                }
            }
            __syncwarp();
            for (int t = lane; t < made; t += 32) {
                int y, f;
                if (replay) { y = pickarr[t]; f = c.first[y]; }
                else {
                    int q = rarr[t];
                    for (int s = t - 1; s >= 0; s--) q += (q >= (int)rarr[s]);
                    int w = 0;
                    if (bwm <= 16) { for (int j = 0; j < bwm - 1; j++) w += (q >= (int)c.cincl[j]); }
                    else {                                   // first word whose inclusive count exceeds q
                        int lo = 0, hi = bwm - 1;
                        while (lo < hi) { const int mid = (lo + hi) >> 1; if (q >= (int)c.cincl[mid]) lo = mid + 1; else hi = mid; }
                        w = lo;
                    }
                    const int before = w ? (int)c.cincl[w - 1] : 0;
                    f = w * 32 + nth_set_bit(c.candw[w], q - before);
                    y = c.la[f];
                }
                // partner degree - 1; dead at 0 (delete_all_pairs_by_id).  Picks are distinct agents: no conflicts.
                const int nd = (int)c.deg[y] - 1;
                c.deg[y] = (int16_t)nd;
                if (nd == 0) atomicAnd(&c.alive[f >> 5], ~(1u << (f & 31)));
                // the row is only reserved here (local indices of the pair); its overlap scan runs after the chain
                const int ri = n_rows + t;
                const uint32_t xy = (uint32_t)x | ((uint32_t)y << 16);
                if (ri >= row_cap) g.ctr->error = MTD_E_CAPACITY;
                else if (ri < stage_cap) stage[ri] = xy;
                else __stcg(&rows[ri].a1, (int32_t)xy);    // beyond the staged range (rare): through the row itself
            }
            if (lane == 0) {                                 // the main agent's degree is spent
                c.deg[x] = (int16_t)(d - spent);
                if (d - spent == 0) { const int f = c.first[x]; atomicAnd(&c.alive[f >> 5], ~(1u << (f & 31))); }
            }
            n_rows += made;
            __syncwarp();
        }
        if (lane == 0) bs.w = n_rows;
    }
    bsync<CW>();
    n_rows = bs.w;
    const int n_out = n_rows < row_cap ? n_rows : row_cap;
    if (tid == 0) g.slot_nrows[slot] = n_out;
    // ---- 7. contact rows of the reserved pairs (contactnetwork.py:98-105, 160-196) ----
    for (int r = tid; r < n_out; r += T) {
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
    bsync<CW>();
}

// Longest-processing-time-first order of one class: counting sort of the class list by interval count, descending,
// into an extra list region (one block).  The order of cells inside a bin is arbitrary; every cell writes only its rows.
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

template <int MM, int CW, int MINB, bool GWS>
__global__ void __launch_bounds__(CW * 32, MINB) k1_block_kernel(DevCtx g, int cls, int list_region, unsigned char *ws,
                                                                  int64_t ws_stride, int mm_rt)
{
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    __shared__ BlockShared bs;
    const Lay L(GWS ? mm_rt : MM);
    unsigned char *sm = GWS ? ws + (int64_t)blockIdx.x * ws_stride : smem_dyn;
    if (g.ctr->error) return;
    const int total = g.ctr->class_cells[cls];
    const int32_t *list = g.class_list + (int64_t)list_region * g.class_stride;
    if (!GWS && threadIdx.x == 0) mbar_init(&bs.mbar, 1);
    if (CW > 1) __syncthreads(); else __syncwarp();
    uint32_t mbar_phase = 0;
    unsigned long long acc_pairs = 0, acc_pot = 0, acc_inc = 0;
    // Slot turnover: the grid holds several times the resident blocks; a block leaves after K1_CELLS_PER_BLOCK cells as long
    // as blocks of its grid are still waiting to start, so the SM slot it held goes to whatever the hardware schedules
    // next -- the short high-priority kernels of the day's other stages (K2/K3, the state chain) included -- instead of
    // staying with one persistent block until the class is exhausted.  Once every block of the grid has started, the
    // blocks that are left drain the class.
    if (threadIdx.x == 0) atomicAdd(&g.ctr->class_started[cls], 1);
    int done = 0;
    while (true) {
        if (threadIdx.x == 0) {
            int w = -1;
            if (done < K1_CELLS_PER_BLOCK || *(volatile int *)&g.ctr->class_started[cls] >= (int)gridDim.x)
                w = atomicAdd(&g.ctr->class_cursor[cls], 1);
            bs.w = w;
        }
        bsync<CW>();
        const int w = bs.w;
        bsync<CW>();                                        // bs.w is reused by the cell
        if (w < 0 || w >= total) break;
        process_cell<CW, GWS>(g, L, sm, bs, mbar_phase, list[w], w + 1 < total ? list[w + 1] : -1, acc_pairs, acc_pot, acc_inc,
                               !GWS && mm_rt < 0, !GWS && g.k1_split && MM <= MTD_K1_SPLIT_MAX_M);
        done++;
    }
    if (threadIdx.x == 0 && (acc_pairs | acc_pot | acc_inc)) {
        atomicAdd(&g.ctr->pairs_evaluated, acc_pairs);
        atomicAdd(&g.ctr->n_potential, acc_pot);
        atomicAdd(&g.ctr->n_incidences, acc_inc);
    }
}

template <int MM, int CW, int MINB>
int launch_block_class(const DevCtx &c, int cls, cudaStream_t st, int sm_count, int lpt_region)
{
    static int per_sm = 0;
    constexpr Lay L(MM);
    if (per_sm == 0) {
        cudaFuncSetAttribute(k1_block_kernel<MM, CW, MINB, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.bytes);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_block_kernel<MM, CW, MINB, false>, CW * 32, L.bytes);
        if (per_sm < 1) per_sm = 1;
    }
    int n = 0;
    if (lpt_region) { k1_lpt_kernel<<<1, 1024, 0, st>>>(c, cls, lpt_region, MM, 512); n++; }
    static const bool skip = [] { const char *e = getenv("MTD_K1_SKIPCHAIN"); return e && *e == '1'; }();
    k1_block_kernel<MM, CW, MINB, false><<<sm_count * per_sm * K1_GRID_ROUNDS, CW * 32, L.bytes, st>>>(c, cls, lpt_region ? lpt_region : cls, nullptr, 0,
                                                                                                    skip ? -MM : MM);
    return n + 1;
}

#define K1_BIG_CW 16
#define K1_BIG_BLOCKS 2

}  // namespace

// bytes of the global workspace the class of cells above 1024 intervals needs (mm_big: its capacity in intervals)
int64_t k1_big_ws_bytes(int mm_big) { const Lay L(mm_big); return (int64_t)K1_BIG_BLOCKS * (int64_t)al16(L.bytes); }

int launch_k1_chain(const DevCtx &c, int cls, int list_region, cudaStream_t st, int sm_count);

// size classes 3..8 (33-64, -128, -192, -320, -1024, above): see size_class() in k0_csr.cu.  lpt[k]: list region holding
// class 3 + k sorted by cell size, largest first (0 = unsorted class list).  With c.k1_split the kernels of classes 3..6
// only prepare their cells and a thread-per-cell chain kernel follows each of them on the same stream (k1_chain.cu).
int launch_k1_blocks(const DevCtx &c, cudaStream_t sa, cudaStream_t sb, int sm_count, unsigned char *big_ws, int mm_big,
                     const int lpt[4], bool only_big, const K1ChainStreams *cs)
{
    int n = 0;
    // the classes with the longest cells first; two streams so that the tail of one class overlaps the next
    if (mm_big > 1024) {
        const Lay L(mm_big);
        k1_block_kernel<1024, K1_BIG_CW, 1, true><<<K1_BIG_BLOCKS, K1_BIG_CW * 32, 0, sa>>>(c, 8, 8, big_ws, (int64_t)al16(L.bytes), mm_big);
        n++;
    }
    if (only_big) return n;
    // the chain kernel of class 3 + k: on its own stream behind the class's prepare kernel (cs), else right behind it
    auto chain = [&](int k, cudaStream_t prep_stream) {
        if (!c.k1_split) return 0;
        cudaStream_t st = prep_stream;
        if (cs) { cudaEventRecord(cs->prep[k], prep_stream); cudaStreamWaitEvent(cs->st[k], cs->prep[k], 0); st = cs->st[k]; }
        const int r = launch_k1_chain(c, 3 + k, lpt[k] ? lpt[k] : 3 + k, st, sm_count);
        if (cs) cudaEventRecord(cs->done[k], st);
        return r;
    };
    static const int sched = [] { const char *e = getenv("MTD_K1_SCHED"); return (e && *e) ? atoi(e) : 1; }();
    if (sched == 1 && cs && c.k1_split) {
        // One stream per size class: ordering helper, prepare kernel and chain kernel of a class follow each other on the
        // class's own stream, and the four classes start together.  A rank of a sharded run holds too few cells to fill the
        // machine, and its K1 stage then lasts as long as the longest dependency chain: lining the classes up two by two on
        // the auxiliary streams put prepare(320) + prepare(128) + chain(128) on one path.
        n += launch_block_class<1024, 8, 1>(c, 7, sa, sm_count, 0);
        for (int k = 3; k >= 0; k--) {
            // prepare kernels of a class follow each other day by day on the class's prepare stream; the chain kernel goes
            // to the chain stream of the day's parity, so that the chains of two consecutive days -- latency-bound, a few
            // resident warps each -- run side by side with the next day's prepare and small-cell kernels
            cudaStream_t st = cs->pst[k] ? cs->pst[k] : cs->st[k];
            cudaStream_t cst = cs->pst[k] ? ((cs->par & 1) ? cs->st2[k] : cs->st[k]) : st;
            cudaStreamWaitEvent(st, cs->fork, 0);
            static const char *nm[4][3] = { { "class 64: fork seen", "class 64: prepared", "class 64: chain done" }, { "class 128: fork seen", "class 128: prepared", "class 128: chain done" },
                                            { "class 192: fork seen", "class 192: prepared", "class 192: chain done" }, { "class 320: fork seen", "class 320: prepared", "class 320: chain done" } };
            k1_tl_mark(nm[k][0], st);
            switch (k) {
            case 3: n += launch_block_class<320, 4, 6>(c, 6, st, sm_count, lpt[3]); break;
            case 2: n += launch_block_class<192, 4, 8>(c, 5, st, sm_count, lpt[2]); break;
            case 1: n += launch_block_class<128, 2, 16>(c, 4, st, sm_count, lpt[1]); break;
            default: n += launch_block_class<64, 2, 16>(c, 3, st, sm_count, lpt[0]); break;
            }
            cudaEventRecord(cs->prep[k], st);             // the class is prepared: its chain kernel is next on this stream
            k1_tl_mark(nm[k][1], st);
            if (cst != st) cudaStreamWaitEvent(cst, cs->prep[k], 0);
            n += launch_k1_chain(c, 3 + k, lpt[k] ? lpt[k] : 3 + k, cst, sm_count);
            cudaEventRecord(cs->done[k], cst);
            k1_tl_mark(nm[k][2], cst);
        }
        return n;
    }
    n += launch_block_class<1024, 8, 1>(c, 7, sa, sm_count, 0);
    n += launch_block_class<320, 4, 6>(c, 6, sa, sm_count, lpt[3]);
    n += chain(3, sa);
    n += launch_block_class<192, 4, 8>(c, 5, sb, sm_count, lpt[2]);
    n += chain(2, sb);
    n += launch_block_class<128, 2, 16>(c, 4, sa, sm_count, lpt[1]);
    n += chain(1, sa);
    n += launch_block_class<64, 2, 16>(c, 3, sb, sm_count, lpt[0]);
    n += chain(0, sb);
    return n;
}
