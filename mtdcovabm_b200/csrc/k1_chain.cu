// K1, split path, kernel 2 of 3 -- the sequential greedy sampling + create_contacts of generate_contact_network
// (contactnetwork.py:283-301, 386-424) with ONE THREAD PER CELL, 32 cells per warp.
//
// Inside a cell the sampling is a chain: every main agent's candidates depend on the picks before it, and a main agent
// draws only a handful of partners, so a warp (let alone a block) per cell leaves almost every lane idle for most of the
// cell's time.  The parallelism that does exist is ACROSS cells.  k1_block.cu therefore only prepares a cell (agents,
// adjacency bit rows, degrees, greedy order) and leaves a record in HBM (K1Rec, common.cuh); here every lane walks the
// main agents of its own cell:
//   candidates = adjacency row of the main agent AND alive          (words in shared memory, lane-interleaved)
//   draw t takes the r_t-th REMAINING candidate in canonical order, r_t = floor(u_t * (nc - t)), Philox keyed
//   (day, cell, main agent, t / 4) -- the same draws as the warp- and block-per-cell kernels, taken one after the other
//   instead of decoded as a Lehmer code;  partner degree - 1, dead at 0 (alive bit cleared = delete_all_pairs_by_id)
// The remaining degrees live in shared memory (u16 per interval position, lane-interleaved), the main agents' records
// are read from HBM in greedy order (sequential per lane, prefetched one record ahead).  Selected pairs are left in the
// cell's scratch rows as (first position of x) | (first position of y) << 16; K2 builds the contact rows from them.
#include "common.cuh"
#include <stdlib.h>

namespace {

#define K1C_WARPS 4

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

// BW: words per adjacency row of the class (even); MMAX: largest cell of the class (intervals)
template <int BW, int MMAX>
__global__ void __launch_bounds__(K1C_WARPS * 32) k1_chain_kernel(DevCtx g, int cls, int list_region)
{
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // per warp: remaining degrees u16[MMAX][32], alive u32[BW][32], candidates u32[BW][32] -- element [i][lane]
    constexpr size_t per_warp = (size_t)MMAX * 64 + (size_t)BW * 128 * 2;
    uint16_t *deg_s = (uint16_t *)(smem + warp * per_warp);
    uint32_t *alive_s = (uint32_t *)(smem + warp * per_warp + (size_t)MMAX * 64);
    uint32_t *cand_s = alive_s + BW * 32;
    if (g.ctr->error) return;
    const int total = g.ctr->class_cells[cls];
    const int32_t *list = g.class_list + (int64_t)list_region * g.class_stride;
    const bool replay = g.day.mode == MTD_MODE_REPLAY;
    int n_miss = 0;
    while (true) {
        int w0 = 0;
        if (lane == 0) w0 = atomicAdd(&g.ctr->class_cursor2[cls], 32);
        w0 = __shfl_sync(FULLMASK, w0, 0);
        if (w0 >= total) break;
        const int idx = w0 + lane;
        if (idx < total) {
            const int slot = list[idx];
            const unsigned char *rec = g.k1_rec + (size_t)64 * (size_t)g.slot_off[slot];
            const int4 hdr = *(const int4 *)rec;
            const int n_act = hdr.x, n = hdr.y, m = hdr.z, bwm = hdr.w;
            int n_rows = 0;
            if (n_act > 0) {
                const K1Rec R(m, n, bwm);
                const int cell = g.slot_cell[slot];
                RowRec *rows = g.scratch_rows + g.slot_rowoff[slot];
                const int row_cap = g.slot_rowbound[slot];
                const uint32_t *mains = (const uint32_t *)(rec + R.o_main);
                const int mw = (int)R.main_words;
                {   // remaining degrees and the alive mask into this lane's shared-memory columns
                    const uint32_t *dsrc = (const uint32_t *)(rec + R.o_deg);
                    for (int p2 = 0; p2 < (m + 1) / 2; p2++) {
                        const uint32_t v = __ldg(dsrc + p2);
                        deg_s[(2 * p2) * 32 + lane] = (uint16_t)(v & 0xffffu);
                        if (2 * p2 + 1 < MMAX) deg_s[(2 * p2 + 1) * 32 + lane] = (uint16_t)(v >> 16);
                    }
                    const uint32_t *asrc = (const uint32_t *)(rec + R.o_alive);
#pragma unroll
                    for (int w = 0; w < BW; w++) alive_s[w * 32 + lane] = w < (int)R.bw2 ? __ldg(asrc + w) : 0u;
                }
                // main record of the current agent in registers, the next one in flight
                uint32_t nx_head = __ldg(mains), nx_agent = __ldg(mains + 1);
                uint32_t nx_row[BW];
#pragma unroll
                for (int w = 0; w < BW; w++) nx_row[w] = w < (int)R.bw2 ? __ldg(mains + 2 + w) : 0u;
                for (int oi = 0; oi < n_act; oi++) {
                    const uint32_t head = nx_head, agx = nx_agent;
                    uint32_t row[BW];
#pragma unroll
                    for (int w = 0; w < BW; w++) row[w] = nx_row[w];
                    if (oi + 1 < n_act) {
                        const uint32_t *nm = mains + (size_t)(oi + 1) * mw;
                        nx_head = __ldg(nm); nx_agent = __ldg(nm + 1);
#pragma unroll
                        for (int w = 0; w < BW; w++) nx_row[w] = w < (int)R.bw2 ? __ldg(nm + 2 + w) : 0u;
                    }
                    const int xpos = (int)(head >> 16);
                    const int dx = deg_s[xpos * 32 + lane];
                    if (dx <= 0) continue;
                    int nc = 0;
#pragma unroll
                    for (int w = 0; w < BW; w++) {           // util.get_all_contacts_ids_by_id: partners with remaining degree > 0
                        const uint32_t cw = row[w] & alive_s[w * 32 + lane];
                        cand_s[w * 32 + lane] = cw;
                        nc += __popc(cw);
                    }
                    if (nc == 0) continue;                     // returned None (contactnetwork.py:294)
                    const int d = dx < nc ? dx : nc;
                    int npick = d, spent = d;
                    long long toff = 0;
                    if (replay) {
                        const long long tr = tape_find_sel(g.tape, cell, (int)agx);
                        if (tr < 0) { n_miss++; continue; }
                        toff = g.tape.sel_off[tr];
                        const int tc = g.tape.sel_off[tr + 1] - (int)toff;
                        if (tc != d) { n_miss++; if (tc < npick) npick = tc; }
                        spent = npick;
                    }
                    uint32_t w4[4] = { 0, 0, 0, 0 };
                    int left = nc;
                    for (int t = 0; t < npick; t++) {
                        int pos = -1;
                        if (replay) {                          // the reference's pick (agent id) -> its first position
                            const int pid = g.tape.sel_picks[toff + t];
                            for (int o2 = 0; o2 < n_act; o2++) {
                                if ((int)__ldg(mains + (size_t)o2 * mw + 1) == pid) { pos = (int)(__ldg(mains + (size_t)o2 * mw) >> 16); break; }
                            }
                            if (pos < 0 || !((cand_s[(pos >> 5) * 32 + lane] >> (pos & 31)) & 1u)) { n_miss++; continue; }
                            cand_s[(pos >> 5) * 32 + lane] &= ~(1u << (pos & 31));
                        } else {
                            if ((t & 3) == 0)
                                philox4x32((uint32_t)g.day.day, (uint32_t)cell, agx, (uint32_t)(t >> 2), g.day.seed, SITE_SELECT, w4);
                            int r = (int)__umulhi(pick4(w4, t & 3), (uint32_t)left);
                            left--;
                            int w = 0;
                            uint32_t cw = cand_s[lane];
                            int cn = __popc(cw);
                            while (r >= cn) { r -= cn; w++; cw = cand_s[w * 32 + lane]; cn = __popc(cw); }
                            const int bit = nth_set_bit(cw, r);
                            cand_s[w * 32 + lane] = cw & ~(1u << bit);
                            pos = w * 32 + bit;
                        }
                        // partner degree - 1; dead at 0 (delete_all_pairs_by_id)
                        const int nd = (int)deg_s[pos * 32 + lane] - 1;
                        deg_s[pos * 32 + lane] = (uint16_t)nd;
                        if (nd == 0) alive_s[(pos >> 5) * 32 + lane] &= ~(1u << (pos & 31));
                        if (n_rows >= row_cap) g.ctr->error = MTD_E_CAPACITY;
                        else __stcg(&rows[n_rows].a1, (int32_t)((uint32_t)xpos | ((uint32_t)pos << 16)));
                        n_rows++;
                    }
                    // the main agent's degree is spent
                    deg_s[xpos * 32 + lane] = (uint16_t)(d - spent);
                    if (d - spent == 0) alive_s[(xpos >> 5) * 32 + lane] &= ~(1u << (xpos & 31));
                }
                g.slot_nrows[slot] = n_rows < row_cap ? n_rows : row_cap;
            }
        }
        __syncwarp();
    }
    if (n_miss) atomicAdd(&g.ctr->tape_miss, n_miss);
}

template <int BW, int MMAX>
int launch_chain(const DevCtx &c, int cls, int list_region, cudaStream_t st, int sm_count)
{
    static int per_sm = 0;
    constexpr size_t smem = ((size_t)MMAX * 64 + (size_t)BW * 128 * 2) * K1C_WARPS;
    if (per_sm == 0) {
        cudaFuncSetAttribute(k1_chain_kernel<BW, MMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_chain_kernel<BW, MMAX>, K1C_WARPS * 32, smem);
        if (per_sm < 1) per_sm = 1;
    }
    k1_chain_kernel<BW, MMAX><<<sm_count * per_sm, K1C_WARPS * 32, smem, st>>>(c, cls, list_region);
    return 1;
}

}  // namespace

// chain kernels of the size classes 3..6 (cells of 33-64, -128, -192, -320 intervals); each on the stream of its class's
// prepare kernel, behind it
int launch_k1_chain(const DevCtx &c, int cls, int list_region, cudaStream_t st, int sm_count)
{
    switch (cls) {
    case 3: return launch_chain<2, 64>(c, cls, list_region, st, sm_count);
    case 4: return launch_chain<4, 128>(c, cls, list_region, st, sm_count);
    case 5: return launch_chain<6, 192>(c, cls, list_region, st, sm_count);
    case 6: return launch_chain<10, 320>(c, cls, list_region, st, sm_count);
    default: return 0;
    }
}
