// K8 -- interval builder: Itinerary.update_cell_agents_timesteps + combine_same_cell_itinerary_entries
// (itinerary.py:1403-1488), the producer of the hot path's input.
//
// Input: one record per itinerary -- its entries {start timestep -> cell} sorted by timestep (CSR ent_off / ent_ts /
// ent_cell; cell -1 = "residence") and the group of agents that share it (CSR grp_off / grp_agent / grp_rescell; a
// local agent is a group of one).  Output: the day's interval stream {agent, cell, start, end} in the reference's
// emission order (itinerary call order, then entry order, then agent order inside the group), i.e. exactly what the
// reference appends to vars_util.cells_agents_timesteps, ready for K0.
//   count : thread per itinerary, walks its entries: runs of consecutive entries in the same cell collapse to the first
//           (:1459-1488), entry i spans [ts_i, ts_{i+1}] with the first starting at 0 and the last ending at 143
//           (:1413-1433), zero-length spans are dropped (:1435) -> intervals x group size
//   scan  : exclusive scan -> output offsets
//   emit  : same walk, writes the records
// Quirk kept: the "-1 = residence" substitution happens inside the agent loop (:1439-1440), so every agent of a group
// gets the residence cell of the FIRST agent of the group.
#include "common.cuh"
#include "scan.cuh"

namespace {

struct ItinCtx {
    int64_t n_itin;
    const int64_t *ent_off; const int16_t *ent_ts; const int32_t *ent_cell;
    const int64_t *grp_off; const int32_t *grp_agent; const int32_t *grp_rescell;
    int32_t *count; const int64_t *out_off;
    int32_t *iv_agent; int32_t *iv_cell; uint8_t *iv_start; uint8_t *iv_end;
    int64_t cap;
    int *error;
};

template <bool EMIT>
__global__ void k8_walk_kernel(ItinCtx c)
{
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < c.n_itin; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t e0 = c.ent_off[t], e1 = c.ent_off[t + 1], g0 = c.grp_off[t], g1 = c.grp_off[t + 1];
        const int gsz = (int)(g1 - g0);
        int64_t out = EMIT ? c.out_off[t] : 0;
        int n_iv = 0;
        // kept entries = first of every run of equal cells; an entry's span ends at the next KEPT entry's timestep
        int64_t k = e0;
        bool first = true;
        while (k < e1) {
            const int cell = c.ent_cell[k];
            int64_t nx = k + 1;
            while (nx < e1 && c.ent_cell[nx] == cell) nx++;                 // collapse the run
            int start = first ? 0 : (int)c.ent_ts[k];
            int end = (nx < e1) ? (int)c.ent_ts[nx] : 143;
            first = false;
            if (start != end) {
                if (EMIT) {
                    const int s = start < 0 ? 0 : start, e = end > 143 ? 143 : end;
                    const int cc = (cell == -1 && gsz > 0) ? c.grp_rescell[g0] : cell;       // quirk: first agent's residence
                    for (int a = 0; a < gsz; a++) {
                        if (out < c.cap) {
                            c.iv_agent[out] = c.grp_agent[g0 + a]; c.iv_cell[out] = cc;
                            c.iv_start[out] = (uint8_t)s; c.iv_end[out] = (uint8_t)(e < 0 ? 0 : e);
                        } else *c.error = MTD_E_CAPACITY;
                        out++;
                    }
                }
                n_iv++;
            }
            k = nx;
        }
        if (!EMIT) c.count[t] = n_iv * gsz;
    }
}

}  // namespace

// count + scan (out_off[n_itin] = total) + emit; returns launches
int launch_k8(int64_t n_itin, const int64_t *ent_off, const int16_t *ent_ts, const int32_t *ent_cell, const int64_t *grp_off,
              const int32_t *grp_agent, const int32_t *grp_rescell, int32_t *count, int64_t *out_off, long long *scan_tmp,
              int32_t *iv_agent, int32_t *iv_cell, uint8_t *iv_start, uint8_t *iv_end, int64_t cap, int *error,
              cudaStream_t st, int sm_count)
{
    ItinCtx c;
    c.n_itin = n_itin; c.ent_off = ent_off; c.ent_ts = ent_ts; c.ent_cell = ent_cell; c.grp_off = grp_off;
    c.grp_agent = grp_agent; c.grp_rescell = grp_rescell; c.count = count; c.out_off = out_off;
    c.iv_agent = iv_agent; c.iv_cell = iv_cell; c.iv_start = iv_start; c.iv_end = iv_end; c.cap = cap; c.error = error;
    const int grid = sm_count * 8;
    k8_walk_kernel<false><<<grid, 256, 0, st>>>(c);
    int n = 1 + device_scan<int32_t, int64_t>(count, out_off, n_itin, nullptr, scan_tmp, st);
    k8_walk_kernel<true><<<grid, 256, 0, st>>>(c);
    return n + 1;
}
