// K0 -- CSR cell -> interval layout of one day's emission stream (replaces the dict-of-lists
// vars_util.cells_agents_timesteps filled at itinerary.py:1446-1452).
//   hist     : per-cell interval count + first emission index (atomics)
//   flags    : record is the first of its cell  -> exclusive scan = slot rank (dict order of the cells)
//   assign   : slot tables, size class lists, per-slot scratch needs
//   scans    : CSR offsets, scratch-row offsets, adjacency-list offsets
//   scatter  : records into their cell segment (unordered; K1 restores list order by emission index)
#include "common.cuh"

#include "scan.cuh"

__device__ __forceinline__ uint32_t order_of(const DevCtx &c, int64_t i) { return c.iv_order ? c.iv_order[i] : (uint32_t)i; }

__global__ void k0_hist_kernel(DevCtx c)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < c.M; i += (int64_t)gridDim.x * blockDim.x) {
        int cell = c.iv_cell[i];
        if (cell < 0 || cell >= c.n_cells) { c.ctr->error = MTD_E_ARG; continue; }
        atomicAdd(&c.cell_count[cell], 1);
        atomicMin(&c.cell_first[cell], order_of(c, i));
    }
}

__global__ void k0_flags_kernel(DevCtx c)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < c.M; i += (int64_t)gridDim.x * blockDim.x) {
        int cell = c.iv_cell[i];
        int f = 0;
        if (cell >= 0 && cell < c.n_cells) f = (c.cell_first[cell] == order_of(c, i));
        c.flag_scan[i] = f;
    }
}

__device__ __forceinline__ int size_class(int m)
{
    return m <= 8 ? 0 : (m <= 16 ? 1 : (m <= 32 ? 2 : (m <= 128 ? 3 : (m <= 320 ? 4 : 5))));
}

__global__ void __launch_bounds__(256) k0_assign_kernel(DevCtx c, const int32_t *flags_excl)
{
    // flags_excl: exclusive scan of the first-of-cell flags, [M+1].  The size-class lists are appended block by block:
    // positions inside the block come from shared-memory counters, one global atomic per class per block iteration
    // (2.2M cells on six global counters would serialise in L2 otherwise).
    __shared__ int s_cnt[MTD_NUM_CLASSES], s_base[MTD_NUM_CLASSES];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t rounds = (c.M + stride - 1) / stride;
    for (int64_t rd = 0; rd < rounds; rd++) {
        const int64_t i = rd * stride + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
        if (threadIdx.x < MTD_NUM_CLASSES) s_cnt[threadIdx.x] = 0;
        __syncthreads();
        int slot = -1, cls = 0, local = 0;
        if (i < c.M) {
            if (i == 0) c.ctr->n_slots = flags_excl[c.M];
            if (flags_excl[i + 1] != flags_excl[i]) {
                slot = flags_excl[i];
                int cell = c.iv_cell[i];
                int m = c.cell_count[cell];
                c.slot_of_cell[cell] = slot;
                c.slot_cell[slot] = cell;
                c.slot_m[slot] = m;
                c.slot_first[slot] = order_of(c, i);
                c.slot_cursor[slot] = 0;
                c.slot_nrows[slot] = 0;
                if (m > MTD_MAX_CELL_INTERVALS) { c.ctr->error = MTD_E_CELL_TOO_LARGE; m = 0; }
                int ct = c.cell_type[cell];
                int indid = (ct == CT_WORKPLACE && c.cell_indid) ? c.cell_indid[cell] : -1;
                long long k = c.P->k_by_act[activity_of_celltype(ct, indid)];
                // rows <= sum(deg)/2, sum(deg) <= max(k*n, S) + n/2 with n <= m; and rows <= m(m-1)/2
                long long b1 = (long long)m * (m - 1) / 2, b2 = (k * m + m / 2 + 2) / 2 + 1;
                c.slot_rowbound[slot] = (int)(b1 < b2 ? b1 : b2);
                c.slot_listneed[slot] = 0;
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

__global__ void k0_check_kernel(DevCtx c)
{
    int n = c.ctr->n_slots;
    c.ctr->row_bound_total = (unsigned long long)c.slot_rowoff[n];
    c.ctr->list_total = (unsigned long long)c.slot_listoff[n];
    if (c.slot_rowoff[n] > c.cap_scratch_rows || c.slot_listoff[n] > c.cap_list) c.ctr->error = MTD_E_CAPACITY;
}

__global__ void k0_scatter_kernel(DevCtx c)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < c.M; i += (int64_t)gridDim.x * blockDim.x) {
        int cell = c.iv_cell[i];
        if (cell < 0 || cell >= c.n_cells) continue;
        int slot = c.slot_of_cell[cell];
        int p = atomicAdd(&c.slot_cursor[slot], 1);
        IvRec r;
        r.agent = c.iv_agent[i];
        r.order = order_of(c, i);
        r.se = (uint32_t)c.iv_start[i] | ((uint32_t)c.iv_end[i] << 8);
        r.pad = 0;
        c.csr[c.slot_off[slot] + p] = r;
    }
}

int launch_k0(const DevCtx &c, long long *scan_tmp, cudaStream_t st, int sm_count)
{
    if (c.M <= 0) return 0;
    int threads = 256;
    int64_t want = (c.M + threads - 1) / threads;
    unsigned grid = (unsigned)(want < (int64_t)sm_count * 16 ? want : (int64_t)sm_count * 16);
    k0_hist_kernel<<<grid, threads, 0, st>>>(c);
    k0_flags_kernel<<<grid, threads, 0, st>>>(c);
    device_scan<int32_t, int32_t>(c.flag_scan, c.flag_scan, c.M, nullptr, scan_tmp, st);
    k0_assign_kernel<<<grid, threads, 0, st>>>(c, c.flag_scan);
    int64_t slots_max = c.M < c.n_cells ? c.M : c.n_cells;
    device_scan<int32_t, int64_t>(c.slot_m, c.slot_off, slots_max, &c.ctr->n_slots, scan_tmp, st);
    device_scan<int32_t, int64_t>(c.slot_rowbound, c.slot_rowoff, slots_max, &c.ctr->n_slots, scan_tmp, st);
    device_scan<int64_t, int64_t>(c.slot_listneed, c.slot_listoff, slots_max, &c.ctr->n_slots, scan_tmp, st);
    k0_check_kernel<<<1, 1, 0, st>>>(c);
    k0_scatter_kernel<<<grid, threads, 0, st>>>(c);
    return 5 + 4 * 3;
}

int launch_scan_nrows(const DevCtx &c, long long *scan_tmp, cudaStream_t st)
{
    if (c.M <= 0) return 0;
    int64_t slots_max = c.M < c.n_cells ? c.M : c.n_cells;
    return device_scan<int32_t, int64_t>(c.slot_nrows, c.slot_denseoff, slots_max, &c.ctr->n_slots, scan_tmp, st);
}
