// K0 -- CSR cell -> interval layout of one day's emission stream (replaces the dict-of-lists
// vars_util.cells_agents_timesteps filled at itinerary.py:1446-1452).
//   hist     : per-cell interval count (the atomic's return value is the record's position inside its cell segment)
//              + first emission index of the cell
//   flags    : "record is the first of its cell" as one bit per record (ballot words) + popcount per word
//   scan     : exclusive scan of the word popcounts (M/32 elements) -> slot rank = dict order of the cells
//   assign   : slot tables, size-class lists, per-slot scratch-row bound
//   scan2    : CSR offsets and scratch-row offsets in ONE fused scan over the slots
//   finish   : per-cell CSR base (so that the scatter needs one gather), capacity check
//   scatter  : records into their cell segment (arrival order; K1 restores list order by emission index)
#include "common.cuh"
#include <stdlib.h>

#include "scan.cuh"

__device__ __forceinline__ uint32_t order_of(const DevCtx &c, int64_t i) { return c.iv_order ? c.iv_order[i] : (uint32_t)i; }

// Intervention coupling of the tiled configurations (the itinerary stage's quarantine /
// hospital / deceased handling, itinerary.py:2283-2528, at day granularity): an agent who is deceased, inside a
// quarantine window or inside a hospitalisation window keeps only the intervals in its residence cell.
__global__ void k0_agent_mask_kernel(DevCtx c)
{
    const int day = c.day.day;
    for (int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; a < c.n_agents; a += (int64_t)gridDim.x * blockDim.x) {
        const int32_t *q = c.quarantine + 3 * a, *h = c.hosp + 3 * a;
        bool m = c.seir[a] == S_DEC;
        const int q0 = q[0], h0 = h[0];
        if (q0 != MTD_NONE && q0 <= day && day <= q[2]) m = true;
        if (h0 != MTD_NONE && h0 <= day && day <= h[2]) m = true;
        c.agent_masked[a] = m;
    }
}

__global__ void k0_hist_kernel(DevCtx c)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < c.M; i += (int64_t)gridDim.x * blockDim.x) {
        int cell = c.iv_cell[i];
        if (cell < 0 || cell >= c.n_cells) { c.ctr->error = MTD_E_ARG; c.iv_pos[i] = 0xffffu; continue; }
        if (c.agent_masked) {
            const int ag = c.iv_agent[i];
            if (c.agent_masked[ag] && cell != c.res_cellid[ag]) { c.iv_pos[i] = 0xffffu; continue; }     // dropped
        }
        const int pos = atomicAdd(&c.cell_count[cell], 1);
        c.iv_pos[i] = (uint16_t)(pos < 0xffff ? pos : 0xffff);         // cells beyond the handle's largest cell are rejected in assign
        atomicMin(&c.cell_first[cell], order_of(c, i));
    }
}

// words[w] bit l: record 32 w + l opens its cell;  cnt[w] = popcount (scanned in place afterwards)
__global__ void k0_flags_kernel(DevCtx c, uint32_t *words, int32_t *cnt)
{
    const int64_t M32 = (c.M + 31) & ~(int64_t)31;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M32; i += (int64_t)gridDim.x * blockDim.x) {
        bool f = false;
        if (i < c.M) {
            int cell = c.iv_cell[i];
            if (cell >= 0 && cell < c.n_cells && c.iv_pos[i] != 0xffffu) f = (c.cell_first[cell] == order_of(c, i));
        }
        const uint32_t w = __ballot_sync(FULLMASK, f);
        if ((threadIdx.x & 31) == 0) { words[i >> 5] = w; cnt[i >> 5] = __popc(w); }
    }
}

__device__ __forceinline__ int size_class(int m)
{
    return m <= 8 ? 0 : (m <= 16 ? 1 : (m <= 32 ? 2 : (m <= 64 ? 3 : (m <= 128 ? 4 : (m <= 192 ? 5 : (m <= 320 ? 6 : (m <= 1024 ? 7 : 8)))))));
}

__global__ void __launch_bounds__(256) k0_assign_kernel(DevCtx c, const uint32_t *words, const int32_t *word_excl)
{
    // word_excl: exclusive scan of the per-word flag counts, [M/32 + 1].  The size-class lists are appended block by
    // block: positions inside the block come from shared-memory counters, one global atomic per class per block
    // iteration (2.2M cells on six global counters would serialise in L2 otherwise).
    __shared__ int s_cnt[MTD_NUM_CLASSES], s_base[MTD_NUM_CLASSES];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t rounds = (c.M + stride - 1) / stride;
    const int lane = threadIdx.x & 31;
    for (int64_t rd = 0; rd < rounds; rd++) {
        const int64_t i = rd * stride + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
        if (threadIdx.x < MTD_NUM_CLASSES) s_cnt[threadIdx.x] = 0;
        __syncthreads();
        int slot = -1, cls = 0, local = 0;
        if (i < c.M) {
            if (i == 0) c.ctr->n_slots = word_excl[(c.M + 31) >> 5];
            const uint32_t w = words[i >> 5];
            if ((w >> lane) & 1u) {
                slot = word_excl[i >> 5] + __popc(w & ((1u << lane) - 1u));
                int cell = c.iv_cell[i];
                int m = c.cell_count[cell];
                c.slot_cell[slot] = cell;
                c.slot_first[slot] = order_of(c, i);
                c.slot_nrows[slot] = 0;
                c.slot_m[slot] = m;                   // the CSR segment keeps its true length; the day is in error
                if (m > c.max_cell_m) { c.ctr->error = MTD_E_CELL_TOO_LARGE; m = 0; }
                int ct = c.cell_type[cell];
                int indid = (ct == CT_WORKPLACE && c.cell_indid) ? c.cell_indid[cell] : -1;
                const int act = activity_of_celltype(ct, indid);
                c.slot_act[slot] = (uint8_t)act;
                long long k = c.P->k_by_act[act];
                // rows <= sum(deg)/2, sum(deg) <= max(k*n, S) + n/2 with n <= m; and rows <= m(m-1)/2
                long long b1 = (long long)m * (m - 1) / 2, b2 = (k * m + m / 2 + 2) / 2 + 1;
                c.slot_rowbound[slot] = (int)(b1 < b2 ? b1 : b2);
                cls = size_class(m);
                local = atomicAdd(&s_cnt[cls], 1);
            }
        }
        __syncthreads();
        if (threadIdx.x < MTD_NUM_CLASSES)
            s_base[threadIdx.x] = s_cnt[threadIdx.x] ? atomicAdd(&c.ctr->class_cells[threadIdx.x], s_cnt[threadIdx.x]) : 0;
        __syncthreads();
        if (slot >= 0) c.class_list[(int64_t)cls * c.class_stride + s_base[cls] + local] = slot;
        __syncthreads();
    }
}

// ---- fused exclusive scan of two int32 arrays over the slots (CSR offsets + scratch-row offsets), same tiling as scan.cuh
__global__ void scan2_reduce_kernel(const int32_t *a, const int32_t *b, int64_t n_max, const int *n_dev, long long *sums)
{
    const int64_t n = min((int64_t)*n_dev, n_max);
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
    long long sa = 0, sb = 0;
    for (int k = threadIdx.x; k < SCAN_TILE; k += SCAN_THREADS) {
        const int64_t i = base + k;
        if (i < n) { sa += a[i]; sb += b[i]; }
    }
    __shared__ long long sh[2][SCAN_THREADS / 32];
    for (int o = 16; o > 0; o >>= 1) { sa += __shfl_down_sync(FULLMASK, sa, o); sb += __shfl_down_sync(FULLMASK, sb, o); }
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = sa; sh[1][threadIdx.x >> 5] = sb; }
    __syncthreads();
    if (threadIdx.x < 2) {
        long long t = 0;
        for (int w = 0; w < SCAN_THREADS / 32; w++) t += sh[threadIdx.x][w];
        sums[(int64_t)threadIdx.x * gridDim.x + blockIdx.x] = t;
    }
}

// two blocks: block y scans sums[y * n_blocks ...] in place
__global__ void scan2_sums_kernel(long long *sums, int64_t n_blocks)
{
    long long *block_sums = sums + (int64_t)blockIdx.x * n_blocks;
    __shared__ long long sh[SCAN_THREADS];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t base = 0; base < n_blocks; base += SCAN_THREADS) {
        const int64_t i = base + threadIdx.x;
        const long long v = (i < n_blocks) ? block_sums[i] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < SCAN_THREADS; o <<= 1) {
            long long t = (threadIdx.x >= o) ? sh[threadIdx.x - o] : 0;
            __syncthreads();
            sh[threadIdx.x] += t;
            __syncthreads();
        }
        const long long incl = sh[threadIdx.x];
        if (i < n_blocks) block_sums[i] = carry + incl - v;
        __syncthreads();
        if (threadIdx.x == SCAN_THREADS - 1) carry += incl;
        __syncthreads();
    }
}

__global__ void scan2_apply_kernel(const int32_t *a, const int32_t *b, int64_t *out_a, int64_t *out_b, int64_t n_max,
                                   const int *n_dev, const long long *sums)
{
    const int64_t n = min((int64_t)*n_dev, n_max);
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
    if (base > n) return;
    constexpr int PER = SCAN_TILE / SCAN_THREADS;
    int va[PER], vb[PER];
    long long sa = 0, sb = 0;
    const int64_t i0 = base + (int64_t)threadIdx.x * PER;
#pragma unroll
    for (int k = 0; k < PER; k++) {
        va[k] = (i0 + k < n) ? a[i0 + k] : 0; vb[k] = (i0 + k < n) ? b[i0 + k] : 0;
        sa += va[k]; sb += vb[k];
    }
    __shared__ long long sh[2][SCAN_THREADS / 32];
    long long ia = sa, ib = sb;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 1; o < 32; o <<= 1) {
        long long ta = __shfl_up_sync(FULLMASK, ia, o), tb = __shfl_up_sync(FULLMASK, ib, o);
        if (lane >= o) { ia += ta; ib += tb; }
    }
    if (lane == 31) { sh[0][w] = ia; sh[1][w] = ib; }
    __syncthreads();
    long long wa = 0, wb = 0;
    for (int k = 0; k < w; k++) { wa += sh[0][k]; wb += sh[1][k]; }
    long long ra = sums[blockIdx.x] + wa + ia - sa, rb = sums[(int64_t)gridDim.x + blockIdx.x] + wb + ib - sb;
#pragma unroll
    for (int k = 0; k < PER; k++) {
        if (i0 + k <= n) { out_a[i0 + k] = ra; out_b[i0 + k] = rb; }      // element n gets the grand totals
        ra += va[k]; rb += vb[k];
    }
}

// per-cell CSR base (one gather in the scatter instead of cell -> slot -> offset) + capacity check
__global__ void k0_finish_kernel(DevCtx c)
{
    const int n = c.ctr->n_slots;
    for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < n; s += gridDim.x * blockDim.x)
        c.cell_base[c.slot_cell[s]] = c.slot_off[s];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        c.ctr->row_bound_total = (unsigned long long)c.slot_rowoff[n];
        if (c.slot_rowoff[n] > c.cap_scratch_rows) c.ctr->error = MTD_E_CAPACITY;
    }
}

__global__ void k0_scatter_kernel(DevCtx c)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < c.M; i += (int64_t)gridDim.x * blockDim.x) {
        const int cell = c.iv_cell[i];
        const uint32_t p = c.iv_pos[i];
        if (cell < 0 || cell >= c.n_cells || p >= (uint32_t)c.max_cell_m) continue;
        IvRec r;
        r.agent = c.iv_agent[i];
        r.order = order_of(c, i);
        r.se = (uint32_t)c.iv_start[i] | ((uint32_t)c.iv_end[i] << 8);
        r.pad = 0;
        c.csr[c.cell_base[cell] + p] = r;
    }
}

__global__ void pack_attr_kernel(DevCtx c, AgentAttr *out)
{
    for (int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; a < c.n_agents; a += (int64_t)gridDim.x * blockDim.x) {
        AgentAttr r; r.soc_rate = c.soc_rate[a]; r.age_bracket = c.age_bracket[a]; r.pad = 0;
        out[a] = r;
    }
}

int launch_pack_attr(const DevCtx &c, AgentAttr *out, cudaStream_t st, int sm_count)
{
    pack_attr_kernel<<<sm_count * 8, 256, 0, st>>>(c, out);
    return 1;
}

int launch_k0(const DevCtx &c, long long *scan_tmp, cudaStream_t st, int sm_count)
{
    if (c.M <= 0) return 0;
    const int threads = 256;
    const int64_t want = (c.M + threads - 1) / threads;
    // blocks per SM's worth of grid: the kernels are grid-stride, so a larger grid means shorter blocks and SM slots that turn
    // over faster for the short kernels of the state chain that run beside K0 (MTD_K0_BLOCKS_PER_SM, default 16)
    static const int bps = [] { const char *e = getenv("MTD_K0_BLOCKS_PER_SM"); const int v = (e && *e) ? atoi(e) : 16; return v < 1 ? 1 : v; }();
    const unsigned grid = (unsigned)(want < (int64_t)sm_count * bps ? want : (int64_t)sm_count * bps);
    const int64_t n_words = (c.M + 31) >> 5;
    int32_t *word_cnt = c.flag_scan;                                  // [n_words + 1], scanned in place
    uint32_t *words = (uint32_t *)(c.flag_scan + n_words + 2);        // [n_words]
    int extra = 0;
    if (c.agent_masked) { k0_agent_mask_kernel<<<sm_count * 8, threads, 0, st>>>(c); extra = 1; }
    k0_hist_kernel<<<grid, threads, 0, st>>>(c);
    k0_flags_kernel<<<grid, threads, 0, st>>>(c, words, word_cnt);
    int n = 2 + device_scan<int32_t, int32_t>(word_cnt, word_cnt, n_words, nullptr, scan_tmp, st);
    k0_assign_kernel<<<grid, threads, 0, st>>>(c, words, word_cnt);
    const int64_t slots_max = c.M < c.n_cells ? c.M : c.n_cells;
    const int64_t nb = (slots_max + 1 + SCAN_TILE - 1) / SCAN_TILE;
    scan2_reduce_kernel<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(c.slot_m, c.slot_rowbound, slots_max, &c.ctr->n_slots, scan_tmp);
    scan2_sums_kernel<<<2, SCAN_THREADS, 0, st>>>(scan_tmp, nb);
    scan2_apply_kernel<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(c.slot_m, c.slot_rowbound, c.slot_off, c.slot_rowoff, slots_max,
                                                             &c.ctr->n_slots, scan_tmp);
    k0_finish_kernel<<<sm_count * 4, threads, 0, st>>>(c);
    k0_scatter_kernel<<<grid, threads, 0, st>>>(c);
    return n + 6 + extra;
}

int launch_scan_nrows(const DevCtx &c, long long *scan_tmp, cudaStream_t st)
{
    if (c.M <= 0) return 0;
    int64_t slots_max = c.M < c.n_cells ? c.M : c.n_cells;
    return device_scan<int32_t, int64_t>(c.slot_nrows, c.slot_denseoff, slots_max, &c.ctr->n_slots, scan_tmp, st);
}
