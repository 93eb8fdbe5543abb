// K1, split path, kernel 2 of 3 -- the sequential greedy sampling + create_contacts of generate_contact_network
// (contactnetwork.py:283-301, 386-424) with a GROUP OF 8 LANES per cell, four cells per warp.
//
// Inside a cell the sampling is a chain: every main agent's candidates depend on the picks before it, and a main agent
// draws only a handful of partners, so a warp (let alone a block) per cell leaves almost every lane idle for most of the
// cell's time.  k1_block.cu therefore only prepares a cell (agents, adjacency bit rows, degrees, greedy order, the first
// Philox block of every main agent) and leaves a record in HBM (K1Rec, common.cuh); here a group of 8 lanes walks the main
// agents of its cell, the four groups of a warp each with their own control flow:
//   candidates = adjacency row of the main agent AND alive (one word per lane, group prefix of the popcounts by shuffles)
//   draw t takes the r_t-th REMAINING candidate in canonical order, r_t = floor(u_t * (nc - t)), Philox keyed
//   (day, cell, main agent, t / 4): a Lehmer code, lane t decodes pick t -- all picks of a main agent in parallel;
//   partner degree - 1, dead at 0 (alive bit cleared = delete_all_pairs_by_id)
// The group's state (remaining degrees, alive / candidate words) lives in shared memory; the main records are read from
// HBM in greedy order, one coalesced 64-byte load per main agent, the next record in flight.  Selected pairs are left in
// the cell's scratch rows as (first position of x) | (first position of y) << 16; K2 builds the contact rows from them.
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

// G lanes per cell (32 / G cells per warp, each group with its own control flow); BW: words per adjacency row of the class
// (even); MMAX: largest cell of the class (intervals); DEG: type of a remaining degree (never above the cell's other agents).
// Per group in shared memory: remaining degrees DEG[MMAX], draws u16[MMAX], alive / candidate / prefix words u32[BW] each,
// the current main record u32[16].
template <int G, int BW, int MMAX, typename DEG>
__global__ void __launch_bounds__(K1C_WARPS * 32) k1_chain_kernel(DevCtx g, int cls, int list_region, int sparse_ok)
{
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int GPW = 32 / G;
    constexpr size_t o_rarr = (MMAX * sizeof(DEG) + 15) & ~(size_t)15, o_alive = o_rarr + ((2 * MMAX + 15) & ~15),
                     o_cand = o_alive + ((4 * BW + 15) & ~15), o_cincl = o_cand + ((4 * BW + 15) & ~15),
                     o_rec = o_cincl + ((4 * BW + 15) & ~15), per_group = o_rec + 64;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gl = lane & (G - 1), gbase = lane & ~(G - 1);
    const uint32_t gmask = (G == 32 ? 0xffffffffu : ((1u << G) - 1u)) << gbase;
    unsigned char *gs = smem + (size_t)(warp * GPW + lane / G) * per_group;
    DEG *deg = (DEG *)gs;
    uint16_t *rarr = (uint16_t *)(gs + o_rarr);
    uint32_t *alive = (uint32_t *)(gs + o_alive), *cand = (uint32_t *)(gs + o_cand), *cincl = (uint32_t *)(gs + o_cincl),
             *recw = (uint32_t *)(gs + o_rec);
    if (g.ctr->error) return;
    const int total = g.ctr->class_cells[cls];
    // Few cells (a rank of a sharded run): one group per warp.  The groups of a warp follow their own control flow, so they
    // take turns at the warp's issue slot, and a cell's chain -- bound by latency -- runs faster alone; with no more cells
    // than warps in the grid every cell still has its group at once.
    if (GPW > 1 && sparse_ok && total <= (int)gridDim.x * K1C_WARPS && lane >= G) return;
    const int32_t *list = g.class_list + (int64_t)list_region * g.class_stride;
    const bool replay = g.day.mode == MTD_MODE_REPLAY;
    int n_miss = 0;
    while (true) {
        int idx = 0;
        if (gl == 0) idx = atomicAdd(&g.ctr->class_cursor2[cls], 1);
        idx = __shfl_sync(gmask, idx, gbase);
        if (idx >= total) break;
        const int slot = list[idx];
        const unsigned char *rec = g.k1_rec + (size_t)MTD_K1_REC_STRIDE * (size_t)g.slot_off[slot];
        const int4 hdr = *(const int4 *)rec;
        const int n_act = hdr.x, n = hdr.y, m = hdr.z, bwm = hdr.w;
        if (n_act <= 0) continue;
        const K1Rec R(m, n, bwm);
        const int cell = g.slot_cell[slot];
        RowRec *rows = g.scratch_rows + g.slot_rowoff[slot];
        const int row_cap = g.slot_rowbound[slot];
        const uint32_t *mains = (const uint32_t *)(rec + R.o_main);
        const int mw = (int)R.main_words;                    // <= 16
        __syncwarp(gmask);                                   // the previous cell's readers of the group's arrays are done
        {
            const uint16_t *dsrc = (const uint16_t *)(rec + R.o_deg);
            for (int p = gl; p < m; p += G) deg[p] = (DEG)dsrc[p];
            const uint32_t *asrc = (const uint32_t *)(rec + R.o_alive);
            for (int w = gl; w < BW; w += G) alive[w] = w < (int)R.bw2 ? asrc[w] : 0u;
        }
        // the main record of the next agent is in flight while the current one is processed: lane gl holds words gl, gl + G
        uint32_t r0 = gl < mw ? __ldg(mains + gl) : 0u, r1 = (G < 16 && gl + G < mw) ? __ldg(mains + gl + G) : 0u;
        int n_rows = 0;
        __syncwarp(gmask);
        for (int oi = 0; oi < n_act; oi++) {
            const uint32_t c0 = r0, c1 = r1;
            if (oi + 1 < n_act) {
                const uint32_t *nm = mains + (size_t)(oi + 1) * mw;
                r0 = gl < mw ? __ldg(nm + gl) : 0u;
                if (G < 16) r1 = gl + G < mw ? __ldg(nm + gl + G) : 0u;
            }
            const int xpos = (int)(__shfl_sync(gmask, c0, gbase) >> 16);
            const int dx = (int)deg[xpos];
            if (dx <= 0) continue;                           // remaining degree 0
            const uint32_t agx = __shfl_sync(gmask, c0, gbase + 1);
            recw[gl] = c0;
            if (G < 16) recw[gl + G] = c1;
            __syncwarp(gmask);
            int nc = 0;
#pragma unroll
            for (int w0 = 0; w0 < BW; w0 += G) {             // candidate words + inclusive popcount prefix
                const int w = w0 + gl;
                const uint32_t cw = w < BW ? (recw[6 + w] & alive[w]) : 0u;     // util.get_all_contacts_ids_by_id
                int sc = __popc(cw);
#pragma unroll
                for (int o = 1; o < G; o <<= 1) { const int t = __shfl_up_sync(gmask, sc, o, G); if (gl >= o) sc += t; }
                if (w < BW) { cand[w] = cw; cincl[w] = (uint32_t)(nc + sc); }
                nc += __shfl_sync(gmask, sc, gbase + G - 1);
            }
            if (nc == 0) continue;                           // returned None (contactnetwork.py:294)
            const int d = dx < nc ? dx : nc;
            int made = d, spent = d;
            __syncwarp(gmask);
            if (replay) {
                // the reference's selected partners (agent ids) -> first positions, validated against the candidate set
                const long long tr = tape_find_sel(g.tape, cell, (int)agx);
                if (tr < 0) { if (gl == 0) n_miss++; continue; }
                const long long toff = g.tape.sel_off[tr];
                const int tc = g.tape.sel_off[tr + 1] - (int)toff;
                int npick = d;
                if (tc != d) { if (gl == 0) n_miss++; if (tc < npick) npick = tc; }
                spent = npick;
                made = 0;
                for (int t = 0; t < npick; t++) {
                    const int pid = g.tape.sel_picks[toff + t];
                    int f = -1;
                    for (int o0 = 0; o0 < n_act; o0 += G) {
                        const int o2 = o0 + gl;
                        const bool hit = o2 < n_act && (int)__ldg(mains + (size_t)o2 * mw + 1) == pid;
                        const uint32_t hb = __ballot_sync(gmask, hit) >> gbase;
                        if (hb) { f = (int)(__ldg(mains + (size_t)(o0 + __ffs(hb) - 1) * mw) >> 16); break; }
                    }
                    const bool ok = f >= 0 && ((cand[f >> 5] >> (f & 31)) & 1u);
                    __syncwarp(gmask);
                    if (!ok) { if (gl == 0) n_miss++; continue; }
                    if (gl == 0) { cand[f >> 5] &= ~(1u << (f & 31)); rarr[made] = (uint16_t)f; }
                    made++;
                    __syncwarp(gmask);
                }
            } else {
                for (int t = gl; t < d; t += G) {
                    uint32_t word;
                    if (t < 4) word = recw[2 + t];           // the agent's first Philox block, drawn by the prepare kernel
                    else {
                        uint32_t w4[4];
                        philox4x32((uint32_t)g.day.day, (uint32_t)cell, agx, (uint32_t)(t >> 2), g.day.seed, SITE_SELECT, w4);
                        word = pick4(w4, t & 3);
                    }
                    rarr[t] = (uint16_t)__umulhi(word, (uint32_t)(nc - t));
                }
            }
            __syncwarp(gmask);
            for (int t = gl; t < made; t += G) {
                int f;
                if (replay) f = rarr[t];
                else {
                    // draw t takes the r_t-th REMAINING candidate: a Lehmer code, decoded by walking the earlier draws backwards
                    int q = rarr[t];
                    for (int s2 = t - 1; s2 >= 0; s2--) q += (q >= (int)rarr[s2]);
                    int w = 0;
#pragma unroll
                    for (int j = 0; j < BW - 1; j++) w += (q >= (int)cincl[j]);
                    const int before = w ? (int)cincl[w - 1] : 0;
                    f = w * 32 + nth_set_bit(cand[w], q - before);
                }
                // partner degree - 1; dead at 0 (delete_all_pairs_by_id).  Picks are distinct agents: no conflicts.
                const int nd = (int)deg[f] - 1;
                deg[f] = (DEG)nd;
                if (nd == 0) atomicAnd(&alive[f >> 5], ~(1u << (f & 31)));
                const int ri = n_rows + t;
                if (ri >= row_cap) g.ctr->error = MTD_E_CAPACITY;
                else __stcg(&rows[ri].a1, (int32_t)((uint32_t)xpos | ((uint32_t)f << 16)));
            }
            if (gl == 0) {                                   // the main agent's degree is spent
                deg[xpos] = (DEG)(d - spent);
                if (d - spent == 0) atomicAnd(&alive[xpos >> 5], ~(1u << (xpos & 31)));
            }
            n_rows += made;
            __syncwarp(gmask);
        }
        if (gl == 0) g.slot_nrows[slot] = n_rows < row_cap ? n_rows : row_cap;
    }
    if (n_miss) atomicAdd(&g.ctr->tape_miss, n_miss);
}

template <int G, int BW, int MMAX, typename DEG>
int launch_chain(const DevCtx &c, int cls, int list_region, cudaStream_t st, int sm_count)
{
    static int per_sm = 0;
    constexpr size_t o_rarr = (MMAX * sizeof(DEG) + 15) & ~(size_t)15, o_alive = o_rarr + ((2 * MMAX + 15) & ~15),
                     per_group = o_alive + 3 * ((4 * BW + 15) & ~15) + 64;
    constexpr size_t smem = per_group * (32 / G) * K1C_WARPS;
    if (per_sm == 0) {
        cudaFuncSetAttribute(k1_chain_kernel<G, BW, MMAX, DEG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_chain_kernel<G, BW, MMAX, DEG>, K1C_WARPS * 32, smem);
        if (per_sm < 1) per_sm = 1;
    }
    static const int sparse_ok = [] { const char *e = getenv("MTD_K1_CHAIN_SPARSE"); return (e && *e == '1') ? 1 : 0; }();      // off by default: faster stand-alone, slower inside the day pipeline (profiles/r02_k1_schedule_ab.txt)
    k1_chain_kernel<G, BW, MMAX, DEG><<<sm_count * per_sm, K1C_WARPS * 32, smem, st>>>(c, cls, list_region, sparse_ok);
    return 1;
}

}  // namespace

// chain kernels of the size classes 3..6 (cells of 33-64, -128, -192, -320 intervals); each on the stream of its class's
// prepare kernel, behind it
int launch_k1_chain(const DevCtx &c, int cls, int list_region, cudaStream_t st, int sm_count)
{
    switch (cls) {
    case 3: return launch_chain<8, 2, 64, uint8_t>(c, cls, list_region, st, sm_count);
    case 4: return launch_chain<8, 4, 128, uint8_t>(c, cls, list_region, st, sm_count);
    case 5: return launch_chain<8, 6, 192, uint8_t>(c, cls, list_region, st, sm_count);
    case 6: return launch_chain<8, 10, 320, uint16_t>(c, cls, list_region, st, sm_count);
    default: return 0;
    }
}
