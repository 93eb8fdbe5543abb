// K1 -- per-cell contact network: ContactNetwork.generate_contact_network (contactnetwork.py:134-315) and the
// row conversion of simulate_contact_network (contactnetwork.py:98-105).
//
// One warp owns one cell at a time (persistent warps fetch cells of their size class from a device cursor).
// Everything the cell needs lives in that warp's slice of shared memory:
//   intervals in list order, local agent index (order of first appearance), per-agent interval lists,
//   the n x n adjacency bit matrix of potential contacts, degrees, candidate buffer.
// The per-agent partner lists in *dict insertion order* (which fixes the candidate order seen by
// util.get_all_contacts_ids_by_id, util.py:274-297) are produced by walking the outer interval loop of the
// reference sequentially and the inner loop in parallel across lanes (ballot + prefix compaction), so the list
// order is the reference's by construction.  Lists live in shared memory for cells of <= 32 intervals and in an
// L2-resident HBM scratch otherwise.
#include "common.cuh"

namespace {

__host__ __device__ constexpr int pow2_ge(int x) { int p = 1; while (p < x) p <<= 1; return p; }
__host__ __device__ constexpr size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

template <int MM>
struct Lay {
    static constexpr int P2 = pow2_ge(MM);
    static constexpr int BW = (MM + 31) / 32;              // bit-matrix words per row
    static constexpr int H = 2 * P2;                       // hash slots
    static constexpr bool LISTS_SMEM = (MM <= 32);
    static constexpr size_t o_u8 = 0;                      // u64[P2]: sort keys / 2 x u32[P2] temps / double[MM] degf
    static constexpr size_t o_agent = o_u8 + align16(sizeof(uint64_t) * P2);        // i32[MM]
    static constexpr size_t o_s = o_agent + align16(sizeof(int32_t) * MM);          // u8[MM]
    static constexpr size_t o_e = o_s + align16(MM);                                // u8[MM]
    static constexpr size_t o_la = o_e + align16(MM);                               // u16[MM]
    static constexpr size_t o_apos = o_la + align16(2 * MM);                        // u16[MM]
    static constexpr size_t o_aoff = o_apos + align16(2 * MM);                      // u16[MM+1]
    static constexpr size_t o_first = o_aoff + align16(2 * (MM + 1));               // u16[MM]
    static constexpr size_t o_pc = o_first + align16(2 * MM);                       // u16[MM]
    static constexpr size_t o_deg = o_pc + align16(2 * MM);                         // i16[MM]
    static constexpr size_t o_ts = o_deg + align16(2 * MM);                         // u16[MM]
    static constexpr size_t o_cand = o_ts + align16(2 * MM);                        // u16[MM]
    static constexpr size_t o_bits = o_cand + align16(2 * MM);                      // u32[MM*BW]
    static constexpr size_t o_hash = o_bits + align16(sizeof(uint32_t) * MM * BW);  // u32[H] (later u16 swap targets)
    static constexpr size_t o_list = o_hash + align16(sizeof(uint32_t) * H);        // u8[MM*MM] if LISTS_SMEM
    static constexpr size_t bytes = o_list + (LISTS_SMEM ? align16((size_t)MM * MM) : 0);
};

struct Cell {
    uint64_t *u8; int32_t *agent; uint8_t *s, *e; uint16_t *la, *apos, *aoff, *first, *pc; int16_t *deg;
    uint16_t *ts, *cand; uint32_t *bits, *hash; uint8_t *slist; uint16_t *glist;
    int bw;          // bit words per row
    int lstride;     // list row stride
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

__device__ __forceinline__ bool bit_get(const Cell &c, int x, int y) { return (c.bits[x * c.bw + (y >> 5)] >> (y & 31)) & 1u; }
__device__ __forceinline__ void bit_set(const Cell &c, int x, int y) { atomicOr(&c.bits[x * c.bw + (y >> 5)], 1u << (y & 31)); }

template <bool SM>
__device__ __forceinline__ int list_get(const Cell &c, int x, int i)
{
    if (SM) return c.slist[x * c.lstride + i];
    return __ldcg(&c.glist[(size_t)x * c.lstride + i]);
}
template <bool SM>
__device__ __forceinline__ void list_put(const Cell &c, int x, int i, int y)
{
    if (SM) c.slist[x * c.lstride + i] = (uint8_t)y;
    else __stcg(&c.glist[(size_t)x * c.lstride + i], (uint16_t)y);
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

__device__ __forceinline__ bool overlap(const Cell &c, int p, int q) { return max(c.s[p], c.s[q]) <= min(c.e[p], c.e[q]); }

// Contact row of the pair {x, y} (local agent indices): key orientation, first overlap's start, last overlap's
// end (contactnetwork.py:101) and the summed overlap length (epidemiology.py:170).
__device__ __forceinline__ void pair_row(const Cell &c, int x, int y, int &a, int &b, int &start, int &end, int &dur)
{
    int fx = 1 << 30, fy = 1 << 30;
    for (int qa = c.aoff[x]; qa < c.aoff[x + 1]; qa++) {
        int pa = c.apos[qa];
        for (int qb = c.aoff[y]; qb < c.aoff[y + 1]; qb++) {
            int pb = c.apos[qb];
            if (overlap(c, pa, pb)) { fx = min(fx, pa); fy = min(fy, pb); }
        }
    }
    a = (fx < fy) ? x : y;          // the key is oriented from the agent whose interval produced the first overlap
    b = (fx < fy) ? y : x;
    bool first = true;
    start = 0; end = 0; dur = 0;
    for (int qa = c.aoff[a]; qa < c.aoff[a + 1]; qa++) {
        int pa = c.apos[qa];
        for (int qb = c.aoff[b]; qb < c.aoff[b + 1]; qb++) {
            int pb = c.apos[qb];
            int lo = max(c.s[pa], c.s[pb]), hi = min(c.e[pa], c.e[pb]);
            if (lo <= hi) {
                if (first) { start = lo; first = false; }
                end = hi;
                dur += hi - lo + 1;
            }
        }
    }
}

__device__ __forceinline__ long long tape_find_sel(const DevTape &T, int cell, int agent)
{
    uint64_t key = ((uint64_t)(uint32_t)cell << 32) | (uint32_t)agent;
    long long lo = 0, hi = T.n_sel;
    while (lo < hi) { long long mid = (lo + hi) >> 1; if (T.sel_key[mid] < key) lo = mid + 1; else hi = mid; }
    return (lo < T.n_sel && T.sel_key[lo] == key) ? lo : -1;
}

template <int MM>
__device__ void process_cell(const DevCtx &g, unsigned char *sm, int slot, int lane, unsigned long long &acc_pairs,
                             unsigned long long &acc_pot)
{
    using L = Lay<MM>;
    constexpr bool SM = L::LISTS_SMEM;
    Cell c;
    c.u8 = (uint64_t *)(sm + L::o_u8); c.agent = (int32_t *)(sm + L::o_agent);
    c.s = sm + L::o_s; c.e = sm + L::o_e; c.la = (uint16_t *)(sm + L::o_la); c.apos = (uint16_t *)(sm + L::o_apos);
    c.aoff = (uint16_t *)(sm + L::o_aoff); c.first = (uint16_t *)(sm + L::o_first); c.pc = (uint16_t *)(sm + L::o_pc);
    c.deg = (int16_t *)(sm + L::o_deg); c.ts = (uint16_t *)(sm + L::o_ts); c.cand = (uint16_t *)(sm + L::o_cand);
    c.bits = (uint32_t *)(sm + L::o_bits); c.hash = (uint32_t *)(sm + L::o_hash);
    c.slist = sm + L::o_list;

    const int m = g.slot_m[slot];
    const int64_t base = g.slot_off[slot];
    const int cell = g.slot_cell[slot];
    const int ctype = g.cell_type[cell];
    const int indid = (ctype == CT_WORKPLACE && g.cell_indid) ? g.cell_indid[cell] : -1;
    const int act = activity_of_celltype(ctype, indid);
    uint32_t *t32a = (uint32_t *)c.u8;            // two u32[P2] temporaries aliasing the sort-key region
    uint32_t *t32b = t32a + L::P2;
    double *degf = (double *)c.u8;

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
        c.u8[p] = k;
    }
    sorted = __all_sync(FULLMASK, sorted);
    __syncwarp();
    if (!sorted) warp_bitonic_sort(c.u8, P2, lane);
    for (int p = lane; p < m; p += 32) {
        IvRec r = g.csr[base + (uint32_t)c.u8[p]];
        c.agent[p] = r.agent; c.s[p] = (uint8_t)(r.se & 0xff); c.e[p] = (uint8_t)((r.se >> 8) & 0xff);
        g.csr_sorted[base + p] = r;
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
    }
    __syncwarp();
    for (int p0 = 0; p0 < m; p0 += 32) {                 // fill the per-agent lists in position order
        int p = p0 + lane;
        int l = (p < m) ? (int)c.la[p] : (0x10000 + lane);
        uint32_t mm = __match_any_sync(FULLMASK, l);
        if (p < m) {
            int rank = __popc(mm & lanemask_lt());
            c.apos[t32a[l] + rank] = (uint16_t)p;
        }
        __syncwarp();
        if (p < m && (__ffs(mm) - 1) == lane) t32a[l] += __popc(mm);
        __syncwarp();
    }
    // per-agent totals, clear the adjacency matrix
    int ts_sum = 0;
    for (int l = lane; l < n; l += 32) {
        int t = 0;
        for (int q = c.aoff[l]; q < c.aoff[l + 1]; q++) { int p = c.apos[q]; int len = (int)c.e[p] - (int)c.s[p] + 1; t += len > 0 ? len : 0; }
        c.ts[l] = (uint16_t)t; ts_sum += t; c.pc[l] = 0; c.deg[l] = 0;
    }
    for (int o = 16; o > 0; o >>= 1) ts_sum += __shfl_xor_sync(FULLMASK, ts_sum, o);
    c.bw = (n + 31) >> 5;
    c.lstride = n > 1 ? n - 1 : 1;
    c.glist = g.lists + g.slot_listoff[slot];
    for (int w = lane; w < n * c.bw; w += 32) c.bits[w] = 0;
    __syncwarp();

    // ---- 3. pair enumeration (contactnetwork.py:160-196): outer loop sequential, inner loop across lanes ----
    for (int i = 0; i < m; i++) {
        const int x = c.la[i];
        const int si = c.s[i], ei = c.e[i];
        int cntx = c.pc[x];
        for (int j0 = 0; j0 < m; j0 += 32) {
            int j = j0 + lane;
            bool f = false;
            int y = 0;
            if (j < m) {
                y = c.la[j];
                f = (y != x) && (max(si, (int)c.s[j]) <= min(ei, (int)c.e[j])) && !bit_get(c, x, y);
                if (f) {                                  // only the first overlapping interval of y creates the key
                    for (int q = c.aoff[y]; q < c.aoff[y + 1]; q++) {
                        int pj = c.apos[q];
                        if (pj >= j) break;
                        if (max(si, (int)c.s[pj]) <= min(ei, (int)c.e[pj])) { f = false; break; }
                    }
                }
            }
            uint32_t bm = __ballot_sync(FULLMASK, f);
            if (f) {
                int r = cntx + __popc(bm & lanemask_lt());
                list_put<SM>(c, x, r, y);                 // key (x, y) appended: x's partner list ...
                int py = c.pc[y];
                list_put<SM>(c, y, py, x);                // ... and y's, in dict insertion order
                c.pc[y] = (uint16_t)(py + 1);
                bit_set(c, x, y); bit_set(c, y, x);
            }
            cntx += __popc(bm);
        }
        if (lane == 0) c.pc[x] = (uint16_t)cntx;
        __syncwarp();
    }
    int pc_sum = 0;
    for (int l = lane; l < n; l += 32) pc_sum += c.pc[l];
    for (int o = 16; o > 0; o >>= 1) pc_sum += __shfl_xor_sync(FULLMASK, pc_sum, o);
    acc_pot += (unsigned)pc_sum / 2;
    if (pc_sum == 0) {                                    // contactnetwork.py:198
        if (lane == 0) g.slot_nrows[slot] = 0;
        __syncwarp();
        return;
    }

    // ---- 4. degrees (contactnetwork.py:210-269) ----
    const mtd_params *P = g.P;
    const long long k = P->k_by_act[act];
    const long long max_degree = k * n;
    const double avg_ts = (double)ts_sum / (double)n;
    const double avg_pc = (double)pc_sum / (double)n;     // every agent has an entry once n >= 2
    const double one_minus_mhd = 1 - g.day.mhd;
    const double log_avg_ts = log(avg_ts), log_avg_pc = log(avg_pc);
    for (int l = lane; l < n; l += 32) {
        double d = 0.0;
        int pc = c.pc[l];
        if (pc > 0) {
            int ag = c.agent[c.first[l]];
            int ab = g.age_bracket[ag];
            if (ab < 0) ab += 16;
            double cmv = P->cm[ab][act];
            double tsm = 1.0, pcm = 1.0;
            double tsv = (double)c.ts[l];
            if (tsv != avg_ts && avg_ts > 1) tsm = log(tsv) / log_avg_ts;
            if ((double)pc != avg_pc && avg_pc > 1) pcm = log((double)pc) / log_avg_pc;
            d = cmv * tsm * pcm * g.soc_rate[ag] * one_minus_mhd;
        }
        degf[l] = d;
    }
    __syncwarp();
    long long deg_sum = 0;
    if (lane == 0) deg_sum = (long long)rint(np_pairwise_sum(degf, n));
    deg_sum = __shfl_sync(FULLMASK, deg_sum, 0);
    for (int l = lane; l < n; l += 32) {
        double d = degf[l];
        long long r;
        if (deg_sum > max_degree) r = (long long)rint((d / (double)deg_sum) * (double)max_degree);
        else r = (long long)rint(d);
        int pc = c.pc[l];
        c.deg[l] = (int16_t)(r > pc ? pc : r);
    }
    __syncwarp();

    // ---- 5. greedy order: sorted(zip(deg, id), reverse=True) (contactnetwork.py:283) ----
    int Pn = 2;
    while (Pn < n) Pn <<= 1;
    for (int l = lane; l < Pn; l += 32) {
        uint64_t key = ~0ull;
        if (l < n) key = ~(((uint64_t)(uint16_t)c.deg[l] << 48) | ((uint64_t)(uint32_t)c.agent[c.first[l]] << 16) | (uint64_t)l);
        c.u8[l] = key;
    }
    __syncwarp();
    warp_bitonic_sort(c.u8, Pn, lane);

    // ---- 6. greedy sampling + create_contacts (contactnetwork.py:287-301, 386-424) ----
    RowRec *rows = g.scratch_rows + g.slot_rowoff[slot];
    const int row_cap = g.slot_rowbound[slot];
    uint16_t *swp = (uint16_t *)c.hash;                   // Fisher-Yates swap targets
    int n_rows = 0;
    for (int oi = 0; oi < n; oi++) {
        const int x = (int)((~c.u8[oi]) & 0xffff);
        int d = c.deg[x];
        if (d <= 0) continue;
        // candidates: partners in insertion order whose remaining degree > 0 (util.py:274-297)
        const int pcx = c.pc[x];
        int nc = 0;
        for (int q0 = 0; q0 < pcx; q0 += 32) {
            int q = q0 + lane;
            int y = 0;
            bool f = false;
            if (q < pcx) { y = list_get<SM>(c, x, q); f = c.deg[y] > 0; }
            uint32_t bm = __ballot_sync(FULLMASK, f);
            if (f) c.cand[nc + __popc(bm & lanemask_lt())] = (uint16_t)y;
            nc += __popc(bm);
        }
        __syncwarp();
        if (nc == 0) continue;                            // get_all_contacts_ids_by_id returned None
        if (d > nc) d = nc;
        const int agx = c.agent[c.first[x]];
        long long toff = -1;
        int npick = d;
        if (g.day.mode == MTD_MODE_REPLAY) {
            long long r = tape_find_sel(g.tape, cell, agx);
            if (r < 0) { if (lane == 0) atomicAdd(&g.ctr->tape_miss, 1); continue; }
            toff = g.tape.sel_off[r];
            int cnt = g.tape.sel_off[r + 1] - (int)toff;
            if (cnt != d) { if (lane == 0) atomicAdd(&g.ctr->tape_miss, 1); if (cnt < npick) npick = cnt; }
        } else {
            // partial Fisher-Yates over the ordered candidate list, Philox keyed by (day, cell, main agent, draw)
            for (int t = lane; t < d; t += 32) {
                uint32_t w[4];
                philox4x32((uint32_t)g.day.day, (uint32_t)cell, (uint32_t)agx, (uint32_t)(t >> 2), g.day.seed, SITE_SELECT, w);
                swp[t] = (uint16_t)(t + (int)__umulhi(w[t & 3], (uint32_t)(nc - t)));
            }
            __syncwarp();
            if (lane == 0) {
                for (int t = 0; t < d; t++) { int r = swp[t]; uint16_t tmp = c.cand[t]; c.cand[t] = c.cand[r]; c.cand[r] = tmp; }
            }
            __syncwarp();
        }
        for (int t0 = 0; t0 < npick; t0 += 32) {
            int t = t0 + lane;
            if (t < npick) {
                int y;
                bool ok = true;
                if (toff >= 0) {
                    int pi = g.tape.sel_picks[toff + t];
                    if (pi < 0 || pi >= nc) { ok = false; atomicAdd(&g.ctr->tape_miss, 1); pi = 0; }
                    y = c.cand[pi];
                } else y = c.cand[t];
                if (ok) {
                    int a, b, st, en, dur;
                    pair_row(c, x, y, a, b, st, en, dur);
                    if (n_rows + t < row_cap) {
                        RowRec r;
                        r.a1 = c.agent[c.first[a]]; r.a2 = c.agent[c.first[b]];
                        r.sed = (uint32_t)st | ((uint32_t)en << 8) | ((uint32_t)dur << 16);
                        r.slot = slot;
                        rows[n_rows + t] = r;
                    } else g.ctr->error = MTD_E_CAPACITY;
                    c.deg[y] -= 1;                        // distinct partners: no two lanes touch the same entry
                }
            }
        }
        if (lane == 0) c.deg[x] = (int16_t)(d - npick);  // main agent's degree is spent
        n_rows += npick;
        __syncwarp();
    }
    if (lane == 0) g.slot_nrows[slot] = n_rows < row_cap ? n_rows : row_cap;
    __syncwarp();
}

template <int MM, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) k1_kernel(DevCtx g, int cls)
{
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char *sm = smem + (size_t)warp * Lay<MM>::bytes;
    if (g.ctr->error) return;
    const int total = g.ctr->class_cells[cls];
    const int32_t *list = g.class_list + (int64_t)cls * g.class_stride;
    unsigned long long acc_pairs = 0, acc_pot = 0;
    while (true) {
        int w = 0;
        if (lane == 0) w = atomicAdd(&g.ctr->class_cursor[cls], 1);
        w = __shfl_sync(FULLMASK, w, 0);
        if (w >= total) break;
        process_cell<MM>(g, sm, list[w], lane, acc_pairs, acc_pot);
    }
    if (lane == 0 && (acc_pairs | acc_pot)) {
        atomicAdd(&g.ctr->pairs_evaluated, acc_pairs);
        atomicAdd(&g.ctr->n_potential, acc_pot);
    }
}

template <int MM, int WARPS>
void launch_class(const DevCtx &c, int cls, cudaStream_t st, int sm_count)
{
    size_t smem = Lay<MM>::bytes * WARPS;
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(k1_kernel<MM, WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        configured = true;
    }
    int per_sm = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_kernel<MM, WARPS>, WARPS * 32, smem);
    if (per_sm < 1) per_sm = 1;
    k1_kernel<MM, WARPS><<<sm_count * per_sm, WARPS * 32, smem, st>>>(c, cls);
}

}  // namespace

int launch_k1(const DevCtx &c, cudaStream_t st, int sm_count)
{
    // big cells first: they are the long poles
    launch_class<1024, 1>(c, 3, st, sm_count);
    launch_class<320, 2>(c, 2, st, sm_count);
    launch_class<128, 4>(c, 1, st, sm_count);
    launch_class<32, 8>(c, 0, st, sm_count);
    return 4;
}
