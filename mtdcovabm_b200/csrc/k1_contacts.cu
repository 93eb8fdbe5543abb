// K1 -- per-cell contact network: ContactNetwork.generate_contact_network (contactnetwork.py:134-315) and the
// row conversion of simulate_contact_network (contactnetwork.py:98-105).
//
// B200 design (no dict, no lists): every agent of a cell gets a 144-bit occupancy mask over the day's 10-minute
// timesteps (union of its intervals).  Two agents are potential contacts iff their masks intersect, which is exactly
// the reference's "some pair of their intervals overlaps" test (contactnetwork.py:426-432, inclusive bounds).  The
// potential-contact graph is therefore an n x n BIT MATRIX built with AND + ballot, potential-contact counts are
// popcounts, and the sequential greedy sampling (contactnetwork.py:283-301, 386-424) works on bit rows:
//   candidates(x) = adj[x] & alive,  alive = {y : remaining degree > 0}
// The reference deletes every pair of an agent whose degree reaches 0 (delete_all_pairs_by_id); with the alive mask
// that deletion is one bit clear.  Selected partners come from the replay tape (agent ids, keyed by cell and main
// agent) or from Philox draws: draw t takes the r_t-th remaining candidate in canonical order (ascending order of
// first appearance in the cell), r_t = floor(u_t * (nc - t)) -- a uniformly random ordered subset, the same
// distribution as np.random.shuffle + np.random.choice(replace=False) (util.py:295, contactnetwork.py:299).
// Only the SELECTED pairs (a few % of the potential ones) need the reference's dict-order details -- key orientation,
// first overlap's start, last overlap's end, summed duration -- and those are recomputed per row from the two
// agents' interval lists (pair_row).
//
// This file:
//   k1_small<G>  cells of <= G intervals (G = 8, 16, 32): G lanes per cell, 32/G cells per warp, one lane per
//                interval and per agent, everything in registers + a few bytes of shared memory per lane.
// Cells of more than 32 intervals: one thread block per cell, k1_block.cu.
#include "common.cuh"
#include <stdlib.h>

#define K1_IMPL_DEFAULT 2      // 0 legacy, 1 block, 2 split (see k1_impl below); split measured fastest on a B200 (K1 3.07 ms serial vs 3.94 legacy / 5.03 block)

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

// Contact row of the pair {x, y}.  ivx / ivy: bit masks (small cells) of the two agents' interval positions, in
// list order; se[]: start | end << 8 per interval position.  Key orientation = the agent whose interval produced the
// first overlap in the reference's loop order (contactnetwork.py:160-196); start = first overlap's start, end = last
// overlap's end (contactnetwork.py:101); dur = summed overlap length (epidemiology.py:170).
__device__ __forceinline__ void pair_row_bits(const uint16_t *se, uint32_t ivx, uint32_t ivy, bool &x_first, int &start,
                                              int &end, int &dur)
{   // one pass: orientation A = x outer / y inner (this loop's order), orientation B = lexicographic (pb, pa)
    if ((ivx & (ivx - 1)) == 0 && (ivy & (ivy - 1)) == 0) {         // one interval each (the common case): one overlap
        const int pa = __ffs(ivx) - 1, pb = __ffs(ivy) - 1;
        start = max((int)(se[pa] & 0xff), (int)(se[pb] & 0xff)); end = min((int)(se[pa] >> 8), (int)(se[pb] >> 8));
        dur = end - start + 1; x_first = pa < pb;
        return;
    }
    int fx = 64, fy = 64;
    int a_start = 0, a_end = 0; bool a_none = true;
    int b_min = 1 << 30, b_max = -1, b_start = 0, b_end = 0;
    dur = 0;
    for (uint32_t bx = ivx; bx; bx &= bx - 1) {
        const int pa = __ffs(bx) - 1; const int sa = se[pa] & 0xff, ea = se[pa] >> 8;
        for (uint32_t by = ivy; by; by &= by - 1) {
            const int pb = __ffs(by) - 1; const int sb = se[pb] & 0xff, eb = se[pb] >> 8;
            const int lo = max(sa, sb), hi = min(ea, eb);
            if (lo <= hi) {
                fx = min(fx, pa); fy = min(fy, pb); dur += hi - lo + 1;
                if (a_none) { a_start = lo; a_none = false; }
                a_end = hi;
                const int kb = (pb << 6) | pa;
                if (kb < b_min) { b_min = kb; b_start = lo; }
                if (kb > b_max) { b_max = kb; b_end = hi; }
            }
        }
    }
    x_first = fx < fy;
    start = x_first ? a_start : b_start;
    end = x_first ? a_end : b_end;
}

// ------------------------------------------------------------------------------------------------ small cells
#define K1S_THREADS 256

// The sub-warp kernels step 32/G cells per warp in lock step: a warp iteration lasts as long as its slowest cell, so
// cells of very different sizes side by side waste lanes (G = 8: four cells of 1..8 intervals per warp).  The class
// lists of the three small classes are therefore counting-sorted by interval count (largest first) before the
// kernels run: neighbours in a warp then hold cells of the same size.  Two passes over the lists (block-aggregated
// histogram, then block-aggregated scatter); the order inside a size bin is arbitrary, every cell writes only its rows.
#define K1B_THREADS 256
__device__ __forceinline__ int small_bin(int cls, int m)
{   // class 0: m in 1..8, class 1: 9..16, class 2: 17..32 -> bin 0 = largest
    const int hi = cls == 0 ? 8 : (cls == 1 ? 16 : 32);
    int b = hi - m;
    return b < 0 ? 0 : (b > 31 ? 31 : b);
}

__global__ void __launch_bounds__(K1B_THREADS) k1_bin_hist_kernel(DevCtx g)
{
    __shared__ int hist[3][32];
    for (int i = threadIdx.x; i < 96; i += K1B_THREADS) (&hist[0][0])[i] = 0;
    __syncthreads();
    for (int cls = 0; cls < 3; cls++) {
        const int total = g.ctr->class_cells[cls];
        const int32_t *list = g.class_list + (int64_t)cls * g.class_stride;
        for (int i = blockIdx.x * K1B_THREADS + threadIdx.x; i < total; i += gridDim.x * K1B_THREADS)
            atomicAdd(&hist[cls][small_bin(cls, g.slot_m[list[i]])], 1);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 96; i += K1B_THREADS) { const int v = (&hist[0][0])[i]; if (v) atomicAdd(&g.k1_bins[i], v); }
}

__global__ void __launch_bounds__(K1B_THREADS) k1_bin_scatter_kernel(DevCtx g, int first_region)
{
    __shared__ int cnt[3][32], base[3][32];
    const int tid = threadIdx.x;
    for (int i = tid; i < 96; i += K1B_THREADS) (&cnt[0][0])[i] = 0;
    __syncthreads();
    // same chunks as the histogram pass: count, reserve one range per bin and block, then place
    for (int cls = 0; cls < 3; cls++) {
        const int total = g.ctr->class_cells[cls];
        const int32_t *list = g.class_list + (int64_t)cls * g.class_stride;
        for (int i = blockIdx.x * K1B_THREADS + tid; i < total; i += gridDim.x * K1B_THREADS)
            atomicAdd(&cnt[cls][small_bin(cls, g.slot_m[list[i]])], 1);
    }
    __syncthreads();
    if (tid < 96) {
        const int cls = tid >> 5, b = tid & 31;
        int off = 0;                                             // start of bin b = sum of the bins before it
        for (int k = 0; k < b; k++) off += g.k1_bins[cls * 32 + k];
        const int c = cnt[cls][b];
        base[cls][b] = off + (c ? atomicAdd(&g.k1_bins[96 + tid], c) : 0);
    }
    __syncthreads();
    for (int cls = 0; cls < 3; cls++) {
        const int total = g.ctr->class_cells[cls];
        const int32_t *list = g.class_list + (int64_t)cls * g.class_stride;
        int32_t *out = g.class_list + (int64_t)(first_region + cls) * g.class_stride;
        for (int i = blockIdx.x * K1B_THREADS + tid; i < total; i += gridDim.x * K1B_THREADS) {
            const int slot = list[i];
            out[atomicAdd(&base[cls][small_bin(cls, g.slot_m[slot])], 1)] = slot;
        }
    }
}


// bin_lo, bin_hi: the kernel takes the cells of the size bins [bin_lo, bin_hi) of the class's size-sorted list (small_bin(): bin
// 0 = the largest cells); bin_hi <= 0 = the whole list.  The class of up to 8 intervals is split this way: G = 8 for its cells of
// 5..8 intervals, G = 4 (eight cells per warp) for those of up to 4, which leave half of an 8-lane group idle.
template <int G>
__global__ void __launch_bounds__(K1S_THREADS, 4) k1_small_kernel(DevCtx g, int cls, int list_region, int bin_lo, int bin_hi)
{
    // per-lane shared slots, indexed [group base + position]
    __shared__ int32_t sh_agent[K1S_THREADS];
    __shared__ uint16_t sh_se[K1S_THREADS];
    __shared__ __align__(16) uint32_t sh_mask[K1S_THREADS][6];     // agent occupancy masks (5 words used)
    __shared__ uint16_t sh_pair[2 * K1S_THREADS];                  // reserved rows of the cell: 2 G pairs per group
    __shared__ __align__(16) uint32_t sh_draw[K1S_THREADS][4];     // first Philox block of the agent on this lane (its first four draws as a main agent)

    const int lane = threadIdx.x & 31;
    const int gl = lane & (G - 1);                  // lane in group
    const int gbase = lane & ~(G - 1);              // first lane of the group inside the warp
    const uint32_t gbits = (G == 32) ? 0xffffffffu : ((1u << G) - 1u);
    const uint32_t gmask = gbits << gbase;
    const int tb = (int)threadIdx.x - gl;           // group base in the block-wide shared arrays
    if (g.ctr->error) return;
    int total = g.ctr->class_cells[cls];            // end of this kernel's range of the list
    int begin = 0;
    if (bin_hi > 0) {
        int acc = 0;
        for (int b = 0; b < 32; b++) {
            if (b == bin_lo) begin = acc;
            if (b == bin_hi) total = acc;
            acc += g.k1_bins[cls * 32 + b];
        }
    }
    const int32_t *list = g.class_list + (int64_t)list_region * g.class_stride;
    const int n_groups = (gridDim.x * K1S_THREADS) / G;
    const mtd_params *P = g.P;
    const double one_minus_mhd = 1 - g.day.mhd;
    unsigned long long acc_pairs = 0, acc_pot = 0, acc_inc = 0;

    // the next cell's descriptor and this lane's record are fetched while the current cell is processed
    int w = begin + (blockIdx.x * K1S_THREADS + (int)threadIdx.x) / G;
    int nx_slot = 0, nx2_slot = 0, nx_m = 0, nx_cell = 0, nx_act = 0; int64_t nx_base = 0;
    IvRec nx_rec; nx_rec.agent = 0; nx_rec.order = 0xffffffffu; nx_rec.se = 0; nx_rec.pad = 0;
    if (w < total) {
        nx_slot = list[w]; nx_m = g.slot_m[nx_slot]; nx_base = g.slot_off[nx_slot]; nx_cell = g.slot_cell[nx_slot];
        nx_act = g.slot_act[nx_slot];
        if (gl < nx_m) nx_rec = g.csr[nx_base + gl];
        if (w + n_groups < total) nx2_slot = list[w + n_groups];       // the slot index runs two cells ahead
    }
    for (; w < total; w += n_groups) {
        const int slot = nx_slot;
        const int m = nx_m;
        const int64_t base = nx_base;
        const int cell = nx_cell;
        IvRec rec = nx_rec;
        const int act = nx_act;
        if (w + n_groups < total) {
            nx_slot = nx2_slot; nx_m = g.slot_m[nx_slot]; nx_base = g.slot_off[nx_slot]; nx_cell = g.slot_cell[nx_slot];
            nx_act = g.slot_act[nx_slot];
            nx_rec.agent = 0; nx_rec.order = 0xffffffffu; nx_rec.se = 0;
            if (gl < nx_m) nx_rec = g.csr[nx_base + gl];
            if (w + 2 * n_groups < total) nx2_slot = list[w + 2 * n_groups];
        }

        if (m < 2) { if (gl == 0) acc_inc += (unsigned)m; continue; }   // a lone interval: no pair, no row (K0 zeroed slot_nrows); with the
                                                    //   size-sorted list these cells sit together at the end
        // ---- 1. my interval; restore list (emission) order by ranking the emission indices ----
        const bool have = gl < m;
        int rank = 0;
#pragma unroll
        for (int j = 0; j < G; j++) {
            uint32_t kj = __shfl_sync(gmask, rec.order, gbase + j);
            rank += (kj < rec.order) || (kj == rec.order && j < gl);
        }
        __syncwarp(gmask);                          // previous cell's readers of the shared slots are done
        if (have) {
            sh_agent[tb + rank] = rec.agent; sh_se[tb + rank] = (uint16_t)(rec.se & 0xffffu);
        }
        __syncwarp(gmask);
        const int agent = have ? sh_agent[tb + gl] : (-1 - gl);
        const uint16_t *se = sh_se + tb;

        // ---- 2. agents = first appearances (contactnetwork.py:161-163); an agent lives on its first interval's lane
        const uint32_t mm = (__match_any_sync(gmask, agent) >> gbase) & gbits;      // my agent's interval positions
        const bool isag = have && ((__ffs(mm) - 1) == gl);
        int ab_pre = 0; double soc_pre = 0.0;                                       // gathered early, used in step 4
        if (isag) { const AgentAttr at = load_attr(g, agent); ab_pre = at.age_bracket; soc_pre = at.soc_rate; }
        const uint32_t agm = (__ballot_sync(gmask, isag) >> gbase) & gbits;         // agent lanes (group-relative)
        const int n = __popc(agm);
        {   // metric: unordered interval pairs of different agents = (m^2 - sum_a cnt_a^2) / 2
            int sq = have ? __popc(mm) : 0;
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) sq += __shfl_xor_sync(gmask, sq, o);
            if (gl == 0) { acc_pairs += (unsigned long long)(m * m - sq) / 2; acc_inc += (unsigned)n; }
        }
        // per-agent occupancy mask and total timesteps (contactnetwork.py:171-174)
        uint32_t mk[5] = { 0, 0, 0, 0, 0 };
        int ts = 0;
        if (isag) {
            for (uint32_t b = mm; b; b &= b - 1) {
                int p = __ffs(b) - 1; int s = se[p] & 0xff, e = se[p] >> 8;
                int len = e - s + 1; ts += len > 0 ? len : 0;
#pragma unroll
                for (int k = 0; k < 5; k++) mk[k] |= range_word(s, e, k);
            }
#pragma unroll
            for (int k = 0; k < 5; k++) sh_mask[tb + gl][k] = mk[k];
        }
        __syncwarp(gmask);

        // ---- 3. potential contacts: bit row per agent (contactnetwork.py:160-196) ----
        uint32_t adj = 0;
        if (isag) {
            for (uint32_t b = agm & ~(1u << gl); b; b &= b - 1) {
                int q = __ffs(b) - 1;
                const uint32_t *o = sh_mask[tb + q];
                uint32_t hit = (mk[0] & o[0]) | (mk[1] & o[1]) | (mk[2] & o[2]) | (mk[3] & o[3]) | (mk[4] & o[4]);
                if (hit) adj |= 1u << q;
            }
        }
        const int pc = __popc(adj);
        int pc_sum = pc, ts_sum = ts;
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) {
            pc_sum += __shfl_xor_sync(gmask, pc_sum, o); ts_sum += __shfl_xor_sync(gmask, ts_sum, o);
        }
        if (gl == 0) acc_pot += (unsigned)pc_sum / 2;
        if (pc_sum == 0) {                                      // contactnetwork.py:198
            if (gl == 0) g.slot_nrows[slot] = 0;
            continue;
        }

        // ---- 4. degrees (contactnetwork.py:210-269) ----
        const long long max_degree = (long long)P->k_by_act[act] * n;
        const double avg_ts = (double)ts_sum / (double)n, avg_pc = (double)pc_sum / (double)n;
        // the two logarithms in ONE pass of the fp64 log sequence: even lanes take avg_ts, odd lanes avg_pc
        const double lg = log((gl & 1) ? avg_pc : avg_ts);
        const double log_avg_ts = __shfl_sync(gmask, lg, gbase), log_avg_pc = __shfl_sync(gmask, lg, gbase + 1);
        double degf = 0.0;
        if (isag && pc > 0) degf = degree_f2(g, P, ab_pre, soc_pre, act, ts, pc, avg_ts, avg_pc, log_avg_ts, log_avg_pc, one_minus_mhd);
        double dsum;
        {   // numpy float64 add.reduce over the agents in order of first appearance (n <= 32 < 128: one block)
            uint32_t b = agm;
            if (n < 8) {
                dsum = -0.0;
                for (; b; b &= b - 1) dsum += __shfl_sync(gmask, degf, gbase + __ffs(b) - 1);
            } else {
                double r[8];
                const int nfull = n - (n % 8);
#pragma unroll
                for (int k = 0; k < 8; k++) { r[k] = __shfl_sync(gmask, degf, gbase + __ffs(b) - 1); b &= b - 1; }
                for (int i = 8; i < nfull; i += 8) {
#pragma unroll
                    for (int k = 0; k < 8; k++) { r[k] += __shfl_sync(gmask, degf, gbase + __ffs(b) - 1); b &= b - 1; }
                }
                dsum = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
                for (; b; b &= b - 1) dsum += __shfl_sync(gmask, degf, gbase + __ffs(b) - 1);
            }
        }
        const long long deg_sum = (long long)rint(dsum);
        int deg = isag ? degree_i(degf, deg_sum, max_degree, pc) : 0;

        // ---- 5. greedy order: sorted(zip(deg, id), reverse=True) (contactnetwork.py:283) ----
        int ord = 0;
        for (uint32_t b = agm; b; b &= b - 1) {
            int q = gbase + __ffs(b) - 1;
            int dq = __shfl_sync(gmask, deg, q), aq = __shfl_sync(gmask, agent, q);
            ord += (dq > deg) || (dq == deg && aq > agent);
        }

        // the first Philox block of every agent that can become a main agent, all agents of the warp's cells in ONE pass of the
        // generator instead of one pass per main agent inside the sequential loop below
        const bool philox_mode = g.day.mode != MTD_MODE_REPLAY;
        if (philox_mode && isag && deg > 0) {
            uint32_t w4[4];
            philox4x32((uint32_t)g.day.day, (uint32_t)cell, (uint32_t)agent, 0u, g.day.seed, SITE_SELECT, w4);
            *(uint4 *)sh_draw[threadIdx.x] = make_uint4(w4[0], w4[1], w4[2], w4[3]);
        }
        __syncwarp(gmask);
        // ---- 6. greedy sampling + create_contacts (contactnetwork.py:287-301, 386-424) ----
        RowRec *rows = g.scratch_rows + g.slot_rowoff[slot];
        const int row_cap = g.slot_rowbound[slot];
        int n_rows = 0;
        for (int r = 0; r < n; r++) {
            const uint32_t xb = __ballot_sync(gmask, isag && ord == r);
            const int x = __ffs(xb) - 1;                           // absolute lane of the main agent
            const int dx = __shfl_sync(gmask, deg, x);
            const uint32_t alive = (__ballot_sync(gmask, isag && deg > 0) >> gbase) & gbits;
            uint32_t cand = __shfl_sync(gmask, adj, x) & alive;     // util.get_all_contacts_ids_by_id
            const int agx = __shfl_sync(gmask, agent, x);
            if (dx <= 0) continue;
            const int nc = __popc(cand);
            if (nc == 0) continue;                                  // returned None (contactnetwork.py:294)
            int d = dx < nc ? dx : nc;
            int npick = d, mypick = -1;
            uint32_t picked = 0;
            if (g.day.mode == MTD_MODE_REPLAY) {
                long long tr = tape_find_sel(g.tape, cell, agx);
                if (tr < 0) { if (gl == 0) atomicAdd(&g.ctr->tape_miss, 1); continue; }
                const int toff = g.tape.sel_off[tr];
                const int cnt = g.tape.sel_off[tr + 1] - toff;
                if (cnt != d) { if (gl == 0) atomicAdd(&g.ctr->tape_miss, 1); if (cnt < npick) npick = cnt; }
                for (int t = 0; t < npick; t++) {
                    const int pid = g.tape.sel_picks[toff + t];     // selected partner's agent id
                    const uint32_t hb = (__ballot_sync(gmask, isag && agent == pid) >> gbase) & gbits & cand;
                    if (!hb) { if (gl == 0) atomicAdd(&g.ctr->tape_miss, 1); continue; }
                    const int bit = __ffs(hb) - 1;
                    cand &= ~(1u << bit); picked |= 1u << bit;
                    if (gl == t) mypick = bit;
                }
            } else {
                // draw t takes the r_t-th REMAINING candidate, r_t = floor(u_t * (nc - t)): a Lehmer code.  Lane t decodes
                // its own pick's rank among the ORIGINAL candidates, q += (q >= r_s) for s = t-1 .. 0, all picks at once.
                int rq = 0;
                if (gl < d) {
                    uint32_t word;
                    if (gl < 4) word = sh_draw[tb + (x - gbase)][gl];
                    else {
                        uint32_t w4[4];
                        philox4x32((uint32_t)g.day.day, (uint32_t)cell, (uint32_t)agx, (uint32_t)(gl >> 2), g.day.seed, SITE_SELECT, w4);
                        word = pick4(w4, gl & 3);
                    }
                    rq = (int)__umulhi(word, (uint32_t)(nc - gl));
                }
                int q = rq;
                for (int s = d - 2; s >= 0; s--) {
                    const int rs = __shfl_sync(gmask, rq, gbase + s);
                    if (s < gl) q += (q >= rs);
                }
                if (gl < d) mypick = nth_set_bit(cand, q);
                picked = __reduce_or_sync(gmask, gl < d ? (1u << mypick) : 0u);
            }
            // rows: lane t reserves the row of pick t (group-relative lanes of the pair); the rows are built after the
            // greedy loop, a lane per row, instead of by d of the G lanes in every iteration
            if (mypick >= 0) {
                const int ri = n_rows + gl;
                const uint32_t xy = (uint32_t)(x - gbase) | ((uint32_t)mypick << 8);
                if (ri >= row_cap) g.ctr->error = MTD_E_CAPACITY;
                else if (ri < 2 * G) sh_pair[2 * tb + ri] = (uint16_t)xy;
                else __stcg(&rows[ri].a1, (int32_t)xy);              // beyond the shared slots (rare): through the row itself
            }
            if ((picked >> gl) & 1u) deg -= 1;                      // partner degree - 1 (dead at 0: alive mask)
            if (lane == x) deg = d - npick;                         // the main agent's degree is spent
            n_rows += npick;
        }
        const int n_out = n_rows < row_cap ? n_rows : row_cap;
        if (gl == 0) g.slot_nrows[slot] = n_out;
        __syncwarp(gmask);
        // ---- 7. contact rows of the reserved pairs (contactnetwork.py:98-105, 160-196) ----
        for (int r0 = 0; r0 < n_out; r0 += G) {
            const int r = r0 + gl;
            const bool v = r < n_out;
            const uint32_t xy = !v ? 0u : (r < 2 * G ? (uint32_t)sh_pair[2 * tb + r] : (uint32_t)__ldcg(&rows[r].a1));
            const int xl = gbase + (int)(xy & 0xffu), yl = gbase + (int)(xy >> 8);
            const uint32_t ivx = __shfl_sync(gmask, mm, xl), ivy = __shfl_sync(gmask, mm, yl);
            const int agx = __shfl_sync(gmask, agent, xl), agy = __shfl_sync(gmask, agent, yl);
            if (v) {
                bool x_first; int st, en, dur;
                pair_row_bits(se, ivx, ivy, x_first, st, en, dur);
                RowRec rr;
                rr.a1 = x_first ? agx : agy; rr.a2 = x_first ? agy : agx;
                rr.sed = (uint32_t)st | ((uint32_t)en << 8) | ((uint32_t)dur << 16);
                rr.slot = slot;
                rows[r] = rr;
            }
        }
    }
    if (gl == 0 && (acc_pairs | acc_pot | acc_inc)) {
        atomicAdd(&g.ctr->pairs_evaluated, acc_pairs);
        atomicAdd(&g.ctr->n_potential, acc_pot);
        atomicAdd(&g.ctr->n_incidences, acc_inc);
    }
}

template <int G>
int launch_small(const DevCtx &c, int cls, cudaStream_t st, int sm_count, int list_region, int bin_lo = 0, int bin_hi = 0)
{
    static int per_sm = 0;
    if (per_sm == 0) {
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_small_kernel<G>, K1S_THREADS, 0);
        if (per_sm < 1) per_sm = 1;
    }
    // Grid of several rounds of the resident blocks: the cells are dealt out grid-stride, so a block of a larger grid
    // simply takes fewer of them and gives its SM slot back sooner -- to the chain kernels of the large classes and the
    // short kernels of the day's other stages, which wait on streams of higher priority.
    // Measured (profiles/r02_k1_schedule_ab.txt): four rounds pay on a full-size day (14M intervals: the state chain gets
    // its SM slots sooner, 4.85 against 5.0 ms per day), one round on the 1.8M-interval shard of an 8-rank run (0.80 against
    // 0.94: with few cells per group the software pipeline of the kernel never fills).
    static const int rounds_env = [] { const char *e = getenv("MTD_K1_SMALL_ROUNDS"); return (e && *e) ? atoi(e) : 0; }();
    const int rounds = rounds_env > 0 ? rounds_env : (c.M >= 8000000 ? 4 : (c.M >= 4000000 ? 2 : 1));
    // MTD_K1_SMALL_PER_SM=n: at most n blocks per SM's worth of grid (4 fill the register file; 3 leave a quarter of it to the
    // short kernels of the state chain, which otherwise wait for a persistent block to end)
    static const int per_sm_cap = [] { const char *e = getenv("MTD_K1_SMALL_PER_SM"); return (e && *e) ? atoi(e) : 0; }();
    // default: on a rank's shard (one round, the blocks stay until the class is done) half of the resident blocks -- the
    // shard's K1 loses nothing (0.715 -> 0.715 ms/day, K1 alone 0.69 -> 0.60: the prepare and chain kernels get the room)
    const int want = per_sm_cap > 0 ? per_sm_cap : (c.M >= 8000000 ? per_sm : (c.M >= 2500000 ? 3 : 2));   // 3.5M intervals (a rank of 4): 1.28 -> 1.23 ms/day with 3
    const int psm = want < per_sm ? want : per_sm;
    k1_small_kernel<G><<<sm_count * psm * rounds, K1S_THREADS, 0, st>>>(c, cls, list_region, bin_lo, bin_hi);
    return 1;
}

}  // namespace

int launch_k1_blocks(const DevCtx &c, cudaStream_t sa, cudaStream_t sb, int sm_count, unsigned char *big_ws, int mm_big,
                     const int lpt[4], bool only_big, const K1ChainStreams *cs);
int launch_k1_legacy(const DevCtx &c, cudaStream_t sa, cudaStream_t sb, int sm_count);

// MTD_K1_IMPL = split  (cells of 33..320 intervals: block per cell prepares, thread per cell samples, K2 builds the rows;
//                        k1_block.cu + k1_chain.cu + k2_rows.cu)
//             | block  (block per cell does everything, k1_block.cu)
//             | legacy (round 1's warp per cell, k1_legacy.cu)
// Cells above 320 (split) resp. 1024 (legacy) intervals go through the block kernels.  All produce the same rows, bit for bit.
int k1_impl()
{
    static const int v = [] { const char *e = getenv("MTD_K1_IMPL"); return !e ? K1_IMPL_DEFAULT : (e[0] == 's' ? 2 : (e[0] == 'b' ? 1 : 0)); }();
    return v;
}
// ---- timeline of one K1 stage (development aid, see common.cuh)
#include <vector>
#include <string>
#include <stdio.h>
static int g_tl_target = [] { const char *e = getenv("MTD_K1_TIMELINE"); return (e && *e) ? atoi(e) : -1; }();
static int g_tl_count = 0;
static bool g_tl_on = false;
static std::vector<std::pair<std::string, cudaEvent_t>> g_tl;
void k1_tl_mark(const char *name, cudaStream_t st)
{
    if (!g_tl_on) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, st);
    g_tl.emplace_back(name, e);
}
static void k1_tl_begin(cudaStream_t st)
{
    if (g_tl_target < 0) return;
    if (g_tl_on && g_tl_count++ == g_tl_target + 1) { k1_tl_mark("NEXT stage begins (main)", st); return; }   // two consecutive stages, one origin
    if (g_tl_on) {                                     // the stage after the recorded ones: print and stop
        cudaDeviceSynchronize();
        for (auto &m : g_tl) { float ms = 0; cudaEventElapsedTime(&ms, g_tl[0].second, m.second); fprintf(stderr, "k1_timeline %8.1f us  %s\n", ms * 1e3, m.first.c_str()); }
        g_tl_on = false; g_tl_target = -1;
        return;
    }
    if (g_tl_count++ == g_tl_target) { g_tl_on = true; k1_tl_mark("stage begins (main)", st); }
}

bool k1_use_legacy() { return k1_impl() == 0; }
bool k1_use_split() { return k1_impl() == 2; }

static int env_int(const char *name, int dflt)
{
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// The block-per-cell kernels of the classes above 32 intervals (k1_block.cu) are bound by the latency of the sequential
// greedy chain of their cells, the sub-warp kernels of the small classes by instruction issue: the two families run side
// by side -- large classes on two auxiliary streams, small classes on the main stream -- and join before K2.
// joined (optional): receives false when the stage's side streams were NOT joined into st -- the caller then makes the
// consumer of the stage wait for ev_join, ev_join2 and cs->done[0..3] itself (see the day overlap in capi.cu).
int launch_k1(const DevCtx &c, cudaStream_t st, cudaStream_t aux, cudaStream_t aux2, cudaEvent_t ev_fork, cudaEvent_t ev_join,
              cudaEvent_t ev_join2, int sm_count, unsigned char *big_ws, int mm_big, const K1ChainStreams *cs, bool *joined)
{
    k1_tl_begin(st);
    static const int overlap = env_int("MTD_K1_OVERLAP", 1);
    static const int lpt = env_int("MTD_K1_LPT", 7);     // bit 0: class 6, bit 1: class 5 longest-first; bit 2: small classes by size
    cudaStream_t sl = st, sl2 = st;
    if (overlap && aux && aux2) {
        cudaEventRecord(ev_fork, st);
        cudaStreamWaitEvent(aux, ev_fork, 0);
        cudaStreamWaitEvent(aux2, ev_fork, 0);
        sl = aux; sl2 = aux2;
    }
    int n;
    const int none[4] = { 0, 0, 0, 0 };
    // list regions of the size-sorted classes: 6 -> +0, 5 -> +1, (small classes +2..+4), 4 -> +5, 3 -> +6
    const int regs[4] = { c.k1_split ? MTD_NUM_CLASSES + 6 : 0, c.k1_split ? MTD_NUM_CLASSES + 5 : 0,
                          (lpt & 2) ? MTD_NUM_CLASSES + 1 : 0, (lpt & 1) ? MTD_NUM_CLASSES : 0 };
    if (k1_use_legacy() && c.k1_bits) {
        n = launch_k1_blocks(c, sl, sl2, sm_count, big_ws, mm_big, none, true, nullptr);
        n += launch_k1_legacy(c, sl, sl2, sm_count);
    } else {
        K1ChainStreams csf;
        if (cs) { csf = *cs; csf.fork = ev_fork; }
        n = launch_k1_blocks(c, sl, sl2, sm_count, big_ws, mm_big, regs, false, (sl != st && cs) ? &csf : nullptr);
    }
    // MTD_K1_SMALL_AFTER_PREP=1: the small-cell kernels start when the large classes are prepared, i.e. together with the
    // chain kernels (which sit on streams of higher priority and take the few SM slots they need first) instead of
    // filling the machine ahead of them and holding it until they end.
    static const int small_after_prep = env_int("MTD_K1_SMALL_AFTER_PREP", 0);
    static const int sched = env_int("MTD_K1_SCHED", 1);
    const bool class_streams = sched == 1 && sl != st && cs && c.k1_split && !(k1_use_legacy() && c.k1_bits);
    const int sorted_small = (lpt & 4) ? MTD_NUM_CLASSES + 2 : 0;      // regions +2, +3, +4 = classes 0, 1, 2 by size
    if (sorted_small) {
        cudaMemsetAsync(c.k1_bins, 0, sizeof(int32_t) * 2 * 96, st);
        k1_bin_hist_kernel<<<sm_count * 2, K1B_THREADS, 0, st>>>(c);
        k1_bin_scatter_kernel<<<sm_count * 2, K1B_THREADS, 0, st>>>(c, sorted_small);
        n += 2;
    }
    if (small_after_prep && class_streams) for (int k = 0; k < 4; k++) cudaStreamWaitEvent(st, cs->prep[k], 0);
    k1_tl_mark("main: before small<16> (bin sort done, prep waits)", st);
    n += launch_small<16>(c, 1, st, sm_count, sorted_small ? sorted_small + 1 : 1);
    k1_tl_mark("main: small<16> done", st);
    // off by default: measured slower (4.61 -> 4.80 ms/day, K1 alone 3.00 -> 3.17: a fourth persistent kernel with its own ramp and
    // tail costs more than the idle lanes of the cells of up to 4 intervals, whose iterations are short anyway); parity-tested
    static const int small4 = env_int("MTD_K1_SMALL4", 0);
    if (sorted_small && small4) {                       // bins of the class: 8 - m, so [0, 4) = 5..8 intervals, [4, 32) = up to 4
        n += launch_small<8>(c, 0, st, sm_count, sorted_small, 0, 4);
        n += launch_small<4>(c, 0, st, sm_count, sorted_small, 4, 32);
    } else n += launch_small<8>(c, 0, st, sm_count, sorted_small ? sorted_small : 0);
    k1_tl_mark("main: small<8> done", st);
    n += launch_small<32>(c, 2, st, sm_count, sorted_small ? sorted_small + 2 : 2);
    k1_tl_mark("main: small<32> done", st);
    if (joined) *joined = true;
    if (sl != st && joined && class_streams) {
        // the caller joins: st carries on with the next day's small cells while this day's chains still run
        cudaEventRecord(ev_join, sl);
        cudaEventRecord(ev_join2, sl2);
        *joined = false;
    } else if (sl != st) {
        cudaEventRecord(ev_join, sl);
        cudaEventRecord(ev_join2, sl2);
        cudaStreamWaitEvent(st, ev_join, 0);
        cudaStreamWaitEvent(st, ev_join2, 0);
        if (cs && c.k1_split && !(k1_use_legacy() && c.k1_bits)) for (int k = 0; k < 4; k++) cudaStreamWaitEvent(st, cs->done[k], 0);
    }
    k1_tl_mark("main: joined", st);
    return n;
}
