// K2 -- dense contact rows in the reference's order (contactnetwork.py:79-105): cells in dict order (slot order),
// rows of a cell in creation order.  K1 wrote each cell's rows into its own scratch segment; here they are packed.
#include "common.cuh"

__device__ __forceinline__ void k2_copy_row(const DevCtx &c, int64_t src, int64_t dst, uint32_t sc)
{
    if (dst < c.cap_rows) {
        const RowRec rr = c.scratch_rows[src];
        c.rows[dst] = rr;
        if (c.rows12) {     // {a1, a2, start | end << 8 | simcelltype << 16}
            uint32_t *o = c.rows12 + 3 * dst;
            o[0] = (uint32_t)rr.a1; o[1] = (uint32_t)rr.a2; o[2] = (rr.sed & 0xffffu) | (sc << 16);
        }
    }
}

// Split K1 (cells of 33..320 intervals): the chain kernel left (first position of x) | (first position of y) << 16 in the
// scratch row; the contact row is built here from the cell's record (K1Rec, common.cuh).  Key orientation = the agent
// whose interval is reached first by the reference's ordered double loop (contactnetwork.py:160-196); start = first
// overlap's start, end = last overlap's end in that orientation (contactnetwork.py:101); dur = summed overlap length
// (epidemiology.py:170).  Same rule as pair_row_lists in k1_block.cu.
__device__ __forceinline__ void k2_build_row(const DevCtx &c, const unsigned char *rec, const K1Rec &R, int slot, int64_t src,
                                             int64_t dst, uint32_t sc)
{
    if (dst >= c.cap_rows) return;
    const uint32_t xy = (uint32_t)__ldcg(&c.scratch_rows[src].a1);
    const uint16_t *la = (const uint16_t *)(rec + R.o_la), *se = (const uint16_t *)(rec + R.o_se),
                   *apos = (const uint16_t *)(rec + R.o_apos), *aoff = (const uint16_t *)(rec + R.o_aoff);
    const int32_t *agent = (const int32_t *)(rec + R.o_agent);
    const int x = la[xy & 0xffffu], y = la[xy >> 16];
    const int xa = aoff[x], xb = aoff[x + 1], ya = aoff[y], yb = aoff[y + 1];
    bool x_first; int start, end, dur = 0;
    if (xb - xa == 1 && yb - ya == 1) {                             // one interval each (the common case): one overlap
        const int pa = apos[xa], pb = apos[ya];
        start = max((int)(se[pa] & 0xff), (int)(se[pb] & 0xff)); end = min((int)(se[pa] >> 8), (int)(se[pb] >> 8));
        dur = end - start + 1; x_first = pa < pb;
    } else {
        int fx = 1 << 30, fy = 1 << 30, a_start = 0, a_end = 0, b_min = 0x7fffffff, b_max = -1, b_start = 0, b_end = 0;
        bool a_none = true;
        for (int qa = xa; qa < xb; qa++) {
            const int pa = apos[qa]; const int sa = se[pa] & 0xff, ea = se[pa] >> 8;
            for (int qb = ya; qb < yb; qb++) {
                const int pb = apos[qb]; const int sb = se[pb] & 0xff, eb = se[pb] >> 8;
                const int lo = max(sa, sb), hi = min(ea, eb);
                if (lo <= hi) {
                    fx = min(fx, pa); fy = min(fy, pb); dur += hi - lo + 1;
                    if (a_none) { a_start = lo; a_none = false; }
                    a_end = hi;
                    const int kb = (pb << 15) | pa;
                    if (kb < b_min) { b_min = kb; b_start = lo; }
                    if (kb > b_max) { b_max = kb; b_end = hi; }
                }
            }
        }
        x_first = fx < fy;
        start = x_first ? a_start : b_start; end = x_first ? a_end : b_end;
    }
    RowRec rr;
    const int agx = agent[x], agy = agent[y];
    rr.a1 = x_first ? agx : agy; rr.a2 = x_first ? agy : agx;
    rr.sed = (uint32_t)start | ((uint32_t)end << 8) | ((uint32_t)dur << 16);
    rr.slot = slot;
    c.rows[dst] = rr;
    if (c.rows12) {
        uint32_t *o = c.rows12 + 3 * dst;
        o[0] = (uint32_t)rr.a1; o[1] = (uint32_t)rr.a2; o[2] = (rr.sed & 0xffffu) | (sc << 16);
    }
}

__global__ void __launch_bounds__(256) k2_compact_kernel(DevCtx c)
{
    const int lane = threadIdx.x & 31;
    const int n_slots = c.ctr->n_slots;
    const int64_t total = c.slot_denseoff[n_slots];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        c.ctr->n_rows = (int)(total < c.cap_rows ? total : c.cap_rows);
        if (total > c.cap_rows) c.ctr->error = MTD_E_CAPACITY;
    }
    const int warps = (gridDim.x * blockDim.x) >> 5;
    const int gwarp = ((int)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    // Cells of more than 32 intervals hold hundreds to thousands of rows: left to the slot loop below, one warp would
    // copy such a cell alone while its 31 other slots wait (the tail of the stage when a rank holds few cells).  They
    // are taken from their size-class lists instead: a warp per cell of class 3 (<= 64 intervals), a block per cell above.
    {
        const int n3 = c.ctr->class_cells[3];
        const int32_t *list = c.class_list + 3 * c.class_stride;
        for (int j = gwarp; j < n3; j += warps) {
            const int s = list[j];
            const int cnt = c.slot_nrows[s];
            if (cnt <= 0) continue;
            const int64_t src = c.slot_rowoff[s], dst = c.slot_denseoff[s];
            const uint32_t sc = c.rows12 ? (uint32_t)simcelltype_of(c.cell_type[c.slot_cell[s]]) : 0u;
            if (c.k1_split) {
                const unsigned char *rec = c.k1_rec + (size_t)MTD_K1_REC_STRIDE * (size_t)c.slot_off[s];
                const int4 hdr = *(const int4 *)rec;
                const K1Rec R(hdr.z, hdr.y, hdr.w);
                for (int r = lane; r < cnt; r += 32) k2_build_row(c, rec, R, s, src + r, dst + r, sc);
            } else
                for (int r = lane; r < cnt; r += 32) k2_copy_row(c, src + r, dst + r, sc);
        }
    }
    for (int cls = 4; cls < MTD_NUM_CLASSES; cls++) {
        const int nl = c.ctr->class_cells[cls];
        const int32_t *list = c.class_list + (int64_t)cls * c.class_stride;
        for (int j = blockIdx.x; j < nl; j += gridDim.x) {
            const int s = list[j];
            const int cnt = c.slot_nrows[s];
            if (cnt <= 0) continue;
            const int64_t src = c.slot_rowoff[s], dst = c.slot_denseoff[s];
            const uint32_t sc = c.rows12 ? (uint32_t)simcelltype_of(c.cell_type[c.slot_cell[s]]) : 0u;
            if (c.k1_split && cls <= 6) {
                const unsigned char *rec = c.k1_rec + (size_t)MTD_K1_REC_STRIDE * (size_t)c.slot_off[s];
                const int4 hdr = *(const int4 *)rec;
                const K1Rec R(hdr.z, hdr.y, hdr.w);
                for (int r = threadIdx.x; r < cnt; r += blockDim.x) k2_build_row(c, rec, R, s, src + r, dst + r, sc);
            } else
                for (int r = threadIdx.x; r < cnt; r += blockDim.x) k2_copy_row(c, src + r, dst + r, sc);
        }
    }
    // Cells of up to 32 intervals hold a handful of rows each (5 on average): the rows of a warp's 32 slots are numbered
    // through (inclusive prefix of the counts in shared memory) and copied 32 at a time, each lane finding its row's slot by
    // a binary search over the prefix -- instead of one slot after the other with a few lanes busy.
    __shared__ int32_t sh_incl[8][32];
    __shared__ int64_t sh_src[8][32], sh_dst[8][32];
    __shared__ uint8_t sh_sct[8][32];
    const int wid = threadIdx.x >> 5;
    for (int s0 = gwarp * 32; s0 < n_slots; s0 += warps * 32) {
        int s = s0 + lane;
        int cnt = 0, sct = 0; int64_t src = 0, dst = 0;
        if (s < n_slots && c.slot_m[s] <= 32) {                 // size classes 0..2 (and the cells of a day in error)
            cnt = c.slot_nrows[s]; src = c.slot_rowoff[s]; dst = c.slot_denseoff[s];
            if (c.rows12 && cnt > 0) sct = simcelltype_of(c.cell_type[c.slot_cell[s]]);
        }
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULLMASK, incl, o); if (lane >= o) incl += t; }
        const int total_w = __shfl_sync(FULLMASK, incl, 31);
        if (total_w == 0) continue;
        __syncwarp();                                           // the previous iteration's readers are done
        sh_incl[wid][lane] = incl; sh_src[wid][lane] = src - (incl - cnt); sh_dst[wid][lane] = dst - (incl - cnt);
        sh_sct[wid][lane] = (uint8_t)sct;
        __syncwarp();
        for (int t = lane; t < total_w; t += 32) {
            int j = 0;                                          // first slot whose inclusive count exceeds t
#pragma unroll
            for (int step = 16; step > 0; step >>= 1) if (sh_incl[wid][j + step - 1] <= t) j += step;
            k2_copy_row(c, sh_src[wid][j] + t, sh_dst[wid][j] + t, (uint32_t)sh_sct[wid][j]);
        }
    }
}

// rows -> the reference's 5-int layout [simcelltype, a1, a2, start, end] (+ duration, cell id)
__global__ void k2_export_kernel(DevCtx c, int32_t *rows5, int32_t *row_dur, int32_t *row_cell, int64_t cap)
{
    int64_t n = c.ctr->n_rows;
    if (n > cap) n = cap;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        RowRec r = c.rows[i];
        int cell = c.slot_cell[r.slot];
        if (rows5) {
            rows5[5 * i + 0] = simcelltype_of(c.cell_type[cell]);
            rows5[5 * i + 1] = r.a1; rows5[5 * i + 2] = r.a2;
            rows5[5 * i + 3] = (int)(r.sed & 0xff); rows5[5 * i + 4] = (int)((r.sed >> 8) & 0xff);
        }
        if (row_dur) row_dur[i] = (int)(r.sed >> 16);
        if (row_cell) row_cell[i] = cell;
    }
}

int launch_k2_compact(const DevCtx &c, cudaStream_t st, int sm_count)
{
    k2_compact_kernel<<<sm_count * 8, 256, 0, st>>>(c);
    return 1;
}

// device -> caller memory (device, or pinned host memory written over PCIe): 16-byte words, count read on the device
__global__ void k2_stream_out_kernel(uint4 *dst, const uint4 *src, const int64_t *n_items, int64_t item_bytes, int64_t cap_items)
{
    int64_t n = *n_items;
    if (n > cap_items) n = cap_items;
    const int64_t bytes = n * item_bytes;
    const int64_t n16 = bytes >> 4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = src[i];
    if (blockIdx.x == 0 && threadIdx.x == 0)            // tail (< 16 bytes; item sizes are multiples of 4)
        for (int64_t b = n16 << 4; b < bytes; b += 4) ((uint32_t *)dst)[b >> 2] = ((const uint32_t *)src)[b >> 2];
}

int launch_stream_out(void *dst, const void *src, const int64_t *n_items, int64_t item_bytes, int64_t cap_items,
                      int blocks, cudaStream_t st)
{
    k2_stream_out_kernel<<<blocks, 256, 0, st>>>((uint4 *)dst, (const uint4 *)src, n_items, item_bytes, cap_items);
    return 1;
}

int launch_export_rows(const DevCtx &c, int32_t *rows5, int32_t *row_dur, int32_t *row_cell, int64_t cap,
                        cudaStream_t st)
{
    k2_export_kernel<<<148 * 4, 256, 0, st>>>(c, rows5, row_dur, row_cell, cap);
    return 1;
}
