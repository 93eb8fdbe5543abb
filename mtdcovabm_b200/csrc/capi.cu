// C ABI of libmtd_b200.so (include/mtd_b200.h): handle, device buffers, entry points.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <vector>
#include "common.cuh"
#include "k5_ctx.h"

// every launcher returns the number of kernels it launched
int launch_k0(const DevCtx &c, long long *scan_tmp, cudaStream_t st, int sm_count);
#define MTD_MAX_PARITIES 3
int launch_k1(const DevCtx &c, cudaStream_t st, cudaStream_t aux, cudaStream_t aux2, cudaEvent_t ev_fork, cudaEvent_t ev_join,
              cudaEvent_t ev_join2, int sm_count, unsigned char *big_ws, int mm_big, const K1ChainStreams *cs, bool *joined);
int64_t k1_big_ws_bytes(int mm_big);
int64_t k1_legacy_bits_words(int sm_count);
bool k1_use_legacy();
bool k1_use_split();
int launch_k2_compact(const DevCtx &c, cudaStream_t st, int sm_count);
int launch_export_rows(const DevCtx &c, int32_t *rows5, int32_t *row_dur, int32_t *row_cell, int64_t cap, cudaStream_t st);
int launch_k3_eval(const DevCtx &c, int32_t *dbuf, int64_t cap, bool win_clean, cudaStream_t st, int sm_count);
int launch_k3_apply(const DevCtx &c, const int32_t *bufs, int n_bufs, int64_t cap, bool resolve,
                    unsigned long long *expo_list, cudaStream_t st, int sm_count);
int launch_k4(const DevCtx &c, int32_t day, const uint8_t *mask, cudaStream_t st, int sm_count);
int launch_set_contacts(const DevCtx &c, int cell, int64_t n, const int32_t *a1, const int32_t *a2, cudaStream_t st);
int launch_k7_schedule(const DevCtx &c, int64_t n_locals, long long k, int32_t *vacc_day, uint8_t *selected, void *scratch,
                       unsigned int *hist, cudaStream_t st, int sm_count);
int launch_k7_replay(const DevCtx &c, int64_t n, const int32_t *ids, const int32_t *ts, int32_t *vacc_day, uint8_t *selected,
                     void *scratch, cudaStream_t st);
int launch_k7_dose(const DevCtx &c, int day, int64_t n_locals, int32_t *vacc_day, cudaStream_t st, int sm_count);
int launch_pack_attr(const DevCtx &c, AgentAttr *out, cudaStream_t st, int sm_count);
int launch_k9_seed(const DevCtx &c, int64_t n, const int32_t *ids, const int32_t *expo_day, int pre_days, const uint8_t *ts_list,
                   int n_ts, const double *asym_mult, const double *draws, int32_t *out, int32_t *status, cudaStream_t st);
int launch_k8(int64_t n_itin, const int64_t *ent_off, const int16_t *ent_ts, const int32_t *ent_cell, const int64_t *grp_off,
              const int32_t *grp_agent, const int32_t *grp_rescell, int32_t *count, int64_t *out_off, long long *scan_tmp,
              int32_t *iv_agent, int32_t *iv_cell, uint8_t *iv_start, uint8_t *iv_end, int64_t cap, int *error,
              cudaStream_t st, int sm_count);
int launch_xchg_push(const DevExchange &x, int par, unsigned long long seq, cudaStream_t st);
int launch_xchg_wait(const DevExchange &x, int par, unsigned long long seq, StateCounters *sctr, cudaStream_t st);
int launch_occupancy(const DevCtx &c, unsigned long long *out6, cudaStream_t st, int sm_count);
int launch_k4_rebuild_next(const DevCtx &c, cudaStream_t st, int sm_count);
int launch_scan_nrows(const DevCtx &c, long long *scan_tmp, cudaStream_t st);
int launch_stream_out(void *dst, const void *src, const int64_t *n_items, int64_t item_bytes, int64_t cap_items,
                      int blocks, cudaStream_t st);

struct mtd_handle_s {
    mtd_config cfg;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaStream_t aux_stream = nullptr, aux2_stream = nullptr;   // large-cell K1 kernels run beside the small-cell ones
    unsigned char *k1_big_ws = nullptr; int k1_mm_big = 0;      // global workspace of the cells above 1024 intervals
    K1ChainStreams k1cs = {}; bool have_k1cs = false;
    unsigned char *k1_rec[MTD_MAX_PARITIES] = {};                              // split K1: per-cell chain records, by day parity (K2 of day d reads
                                                                //   them while the K1 of day d+1 already writes)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_join2 = nullptr;
    // The state chain -- apply of day d's exposure records (after the peer exchange), then K4 of day d+1 -- runs on its
    // own stream beside K0/K1/K2 of day d+1, which read no per-agent state; K3 of day d+1 joins the two.
    cudaStream_t state_stream = nullptr;
    cudaEvent_t ev_state = nullptr, ev_main = nullptr;
    // K0 of day d+1 reads only the interval stream and static cell attributes, and it is bound by L2 atomics and a random
    // scatter while K1 is bound by instruction issue: it runs on a stream of its own beside K1..K3 of day d.  Everything
    // K0 produces (and the day's counters) is double-buffered by day parity; DevCtx is passed by value at launch, so
    // swapping the pointers in h->c at the start of a day is all the later stages need.
    struct K0Set { int32_t *cell_count; uint32_t *cell_first; int64_t *cell_base; uint16_t *iv_pos; int32_t *flag_scan;
                   int32_t *slot_cell, *slot_m; uint32_t *slot_first; uint8_t *slot_act; int64_t *slot_off, *slot_rowoff;
                   int32_t *slot_rowbound, *slot_nrows; int64_t *slot_denseoff; int32_t *class_list; IvRec *csr;
                   DevCounters *ctr; RowRec *scratch_rows; } k0set[MTD_MAX_PARITIES] = {};
    int n_par = 2;                               // sets of the K0 / K1 products in flight (MTD_PARITIES = 2 | 3)
    cudaStream_t k1_join_stream = nullptr; cudaEvent_t ev_k1_small = nullptr; bool k1_days_overlap = false;   // see the day overlap of K1 in mtd_simulate_contact_network
    cudaStream_t k1_stream2 = nullptr; int32_t *k1_bins_base = nullptr;   // K1 stages of consecutive days alternate between k1_stream and k1_stream2 (day overlap)
    cudaStream_t k0_stream = nullptr, k1_stream = nullptr;   // K1 of day d+1 likewise runs beside K2/K3 of day d (its scratch rows are per parity too)
    cudaEvent_t ev_k1_done = nullptr;
    bool last_pipe = false;                      // the previous day ran pipelined
    bool k1_pipe = true;                         // MTD_K1_PIPE=0 keeps K1 on the main stream
    cudaEvent_t ev_k0_ready = nullptr, ev_main_mark[MTD_MAX_PARITIES] = {};
    int32_t *n_succ_buf = nullptr;               // [2] by day parity (DevCtx::n_succ_pub)
    long long *scan_tmp_k0 = nullptr;
    bool k0_pipe = true;                         // MTD_K0_PIPE=0 keeps K0 on the main stream
    bool k0_pipe_staged = false;
    bool k0_device_ready = false;                // mtd_set_k0_overlap: MTD_MEM_DEVICE inputs are complete at call time
    int64_t k0_seq = 0;
    bool state_dirty = false;                    // the state stream has work the main stream has not waited for
    bool main_dirty = false;                     // the main stream touched per-agent state since the state stream last waited
    // peer exchange (NVLink peer memory, one slice per source rank and day parity)
    DevExchange x;
    bool xchg_on = false;
    void *xchg_alloc = nullptr;
    std::vector<void *> xchg_peers;              // cudaIpcOpenMemHandle mappings
    unsigned long long xchg_seq = 0;
    // caller-supplied contacts (mtd_set_contacts), device copies
    int32_t *sc_a1 = nullptr, *sc_a2 = nullptr; int64_t sc_cap_rows = 0; int64_t *sc_off = nullptr; int64_t sc_cap_off = 0;
    uint8_t *sc_se = nullptr; int64_t sc_cap_se = 0;
    uint8_t *agent_masked = nullptr; bool mask_on = false;
    int32_t *k6_agent = nullptr; double *k6_u = nullptr; int64_t k6_n = 0;   // replay tape of K6 (mtd_set_k6_tape)
    TraceCtx *trace_ctx = nullptr; bool trace_scan_pending = false;   // cell-sharded K5: context between scan and finish
    uint8_t *seed_buf = nullptr; int64_t seed_cap = 0;        // staging of mtd_seed_exposures
    // K8 interval builder: device copies of the itinerary CSR + scratch; the built stream lives in d_iv_*
    int64_t *k8_ent_off = nullptr, *k8_grp_off = nullptr, *k8_out_off = nullptr; int16_t *k8_ent_ts = nullptr;
    int32_t *k8_ent_cell = nullptr, *k8_grp_agent = nullptr, *k8_grp_res = nullptr, *k8_count = nullptr; long long *k8_scan = nullptr;
    int64_t k8_cap[9] = {}; int64_t built_n = -1;
    // K7 vaccination (allocated by mtd_enable_vaccination)
    int32_t *vacc_day = nullptr; uint8_t *vacc_selected = nullptr; void *vacc_scratch = nullptr; unsigned int *vacc_hist = nullptr;
    int64_t vacc_n_locals = 0; int32_t *vacc_ids = nullptr, *vacc_ts = nullptr; int64_t vacc_cap_ids = 0, vacc_cap_ts = 0;
    int sm_count = 148;
    std::string err;
    std::vector<void *> allocs;
    DevCtx c;
    // staging buffers for host-resident day inputs
    int32_t *d_iv_agent = nullptr, *d_iv_cell = nullptr; uint8_t *d_iv_start = nullptr, *d_iv_end = nullptr;
    uint32_t *d_iv_order = nullptr;
    uint8_t *d_mask = nullptr;
    long long *scan_tmp = nullptr;
    unsigned long long *expo_list = nullptr;     // [cap_expo][2] (key, agent)
    unsigned long long *occ = nullptr;
    mtd_params *d_params = nullptr;
    int32_t *delta_own = nullptr;
    // replay tape (device copies)
    std::vector<void *> tape_allocs;
    bool have_params = false, have_cells = false, have_static = false, day_ready = false;
    // K5 (allocated on first use)
    int64_t n_res_cells = 0, n_res_members = 0;
    int64_t *res_off = nullptr; int32_t *res_nres = nullptr, *res_members = nullptr, *trace_sel = nullptr;
    int32_t *res_first = nullptr, *res_k = nullptr;
    uint8_t *is_traced = nullptr, *primary = nullptr, *secondary = nullptr, *d_epi_present = nullptr;
    int32_t *trace_order = nullptr, *trace_cursor = nullptr, *sec_owner = nullptr, *trace_scalars = nullptr;
    int32_t *trace_flagscan = nullptr; long long *trace_scan_tmp = nullptr;
    int32_t *d_traced = nullptr; int64_t cap_traced = 0; int32_t *d_rows3 = nullptr; int64_t cap_rows3 = 0;
    std::vector<void *> trace_tape_allocs;
    // pipelined boundary: staged uploads (copy stream) and streamed exports (side stream), double-buffered
    cudaStream_t copy_stream = nullptr, export_stream = nullptr, bulk_stream = nullptr;
    int32_t *st_agent[2] = {}, *st_cell[2] = {}; uint8_t *st_start[2] = {}, *st_end[2] = {}; uint32_t *st_order[2] = {};
    int64_t st_n[2] = {}; bool st_has_order[2] = {};
    int st_next = 0, st_ready = -1;                 // slot the next prefetch fills / slot holding the prefetched day
    cudaEvent_t ev_staged[2] = {}, ev_k0_done = nullptr, ev_snap = nullptr, ev_exported[2] = {};
    bool export_on = false;
    bool win_clean = false;                         // the winner table holds no entries (cleared after the last apply)
    uint32_t *rows12_buf[2] = {};                   // packed rows of the last two days
    int64_t *snap[2] = {};                          // device copy of mtd_day_result per export slot
    unsigned long long *expo_snap[2] = {};
    int64_t day_seq = 0, export_ticket = 0;         // days computed / exports enqueued
    bool exported_pending[2] = {};
    int ticket_slot[2] = {};
    // host exports: the result record lands first (ev_res); the bulk copies are issued with exact sizes once it has
    struct HostExport { bool unissued = false; void *rows = nullptr; int64_t cap_rows = 0; void *kv = nullptr;
                        int64_t cap_exposed = 0; mtd_day_result *result = nullptr; } hx[2];
    cudaEvent_t ev_res[2] = {};
    // measurement hooks
    bool profiling = false;
    // ring of event pairs per stage: every execution since the last mtd_get_stage_ms is timed (up to EV_RING)
    static constexpr int EV_RING = 512;
    std::vector<cudaEvent_t> ev[MTD_N_STAGES];      // [2 * EV_RING], created on first use
    int ev_n[MTD_N_STAGES] = {};                    // executions recorded since the last query
    int64_t launches[MTD_N_STAGES] = {};
};

struct StageScope {     // records the stage's start/end events on the stage's stream when profiling is on
    mtd_handle h; int stage; cudaStream_t st;
    StageScope(mtd_handle h_, int s, cudaStream_t st_ = nullptr) : h(h_), stage(s), st(st_ ? st_ : h_->stream)
    { if (on()) cudaEventRecord(h->ev[stage][2 * h->ev_n[stage]], st); }
    ~StageScope() { if (on()) { cudaEventRecord(h->ev[stage][2 * h->ev_n[stage] + 1], st); h->ev_n[stage]++; } }
    bool on() const { return h->profiling && h->ev_n[stage] < mtd_handle_s::EV_RING; }
    void add(int n) { h->launches[stage] += n; }
};

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    h->err = std::string(#call) + ": " + cudaGetErrorString(e_); return MTD_E_CUDA; } } while (0)

template <typename T>
static int dalloc(mtd_handle h, T **p, size_t n)
{
    void *q = nullptr;
    cudaError_t e = cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T));
    if (e != cudaSuccess) { h->err = std::string("cudaMalloc: ") + cudaGetErrorString(e); return MTD_E_CUDA; }
    h->allocs.push_back(q);
    *p = (T *)q;
    return 0;
}
#define DA(ptr, n) do { int r_ = dalloc(h, &(ptr), (size_t)(n)); if (r_) return r_; } while (0)

// main stream waits for everything enqueued on the state stream so far
static int main_after_state(mtd_handle h)
{
    if (!h->state_dirty) return MTD_OK;
    CK(cudaEventRecord(h->ev_state, h->state_stream));
    CK(cudaStreamWaitEvent(h->stream, h->ev_state, 0));
    h->state_dirty = false;
    return MTD_OK;
}
// state stream waits for everything enqueued on the main stream so far
static int state_after_main(mtd_handle h)
{
    CK(cudaEventRecord(h->ev_main, h->stream));
    CK(cudaStreamWaitEvent(h->state_stream, h->ev_main, 0));
    h->main_dirty = false;
    return MTD_OK;
}
// entry of every call that reads or writes per-agent state on the main stream outside the day step
static int touch_state_on_main(mtd_handle h)
{
    int r = main_after_state(h);
    h->main_dirty = true;
    return r;
}
#define TOUCH() do { int r_ = touch_state_on_main(h); if (r_) return r_; } while (0)

extern "C" int mtd_abi_version(void) { return 3; }

__global__ void debug_log_kernel(int64_t n, const double *x, double *out)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = log(x[i]);
}

// test hook (tests/test_device_log.py): CUDA's fp64 log on the caller's values, as K1 computes log(avg_ts), log(avg_pc)
extern "C" int mtd_debug_log(mtd_handle h, int64_t n, const double *x, double *out)
{
    if (!h || n < 0 || (n && (!x || !out))) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    if (n == 0) return MTD_OK;
    double *dx = nullptr, *dout = nullptr;
    CK(cudaMalloc((void **)&dx, 8 * (size_t)n)); 
    if (cudaMalloc((void **)&dout, 8 * (size_t)n) != cudaSuccess) { cudaFree(dx); h->err = "mtd_debug_log: cudaMalloc"; return MTD_E_CUDA; }
    cudaMemcpyAsync(dx, x, 8 * (size_t)n, cudaMemcpyHostToDevice, h->stream);
    debug_log_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(n, dx, dout);
    cudaMemcpyAsync(out, dout, 8 * (size_t)n, cudaMemcpyDeviceToHost, h->stream);
    cudaError_t e = cudaStreamSynchronize(h->stream);
    cudaFree(dx); cudaFree(dout);
    if (e != cudaSuccess) { h->err = std::string("mtd_debug_log: ") + cudaGetErrorString(e); return MTD_E_CUDA; }
    return MTD_OK;
}

extern "C" const char *mtd_last_error(mtd_handle h) { return h ? h->err.c_str() : "null handle"; }

static int create_impl(mtd_handle h)
{
    const mtd_config &cfg = h->cfg;
    CK(cudaSetDevice(cfg.device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, cfg.device));
    h->sm_count = prop.multiProcessorCount;
    if (cfg.stream) h->stream = (cudaStream_t)cfg.stream;
    else { CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)); h->own_stream = true; }
    CK(cudaStreamCreateWithFlags(&h->aux_stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&h->aux2_stream, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&h->ev_join2, cudaEventDisableTiming));
    {   // the state chain (apply, K4) is short and on the path to the next K3: highest priority, like the caller's main stream
        int plo = 0, phi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&plo, &phi));
        const char *pe = getenv("MTD_MAIN_PRIO");
        CK(cudaStreamCreateWithPriority(&h->state_stream, cudaStreamNonBlocking, (pe && *pe == '0') ? plo : phi));
    }
    CK(cudaEventCreateWithFlags(&h->ev_state, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&h->ev_main, cudaEventDisableTiming));
    memset(&h->x, 0, sizeof(h->x));
    CK(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
    DevCtx &c = h->c;
    memset(&c, 0, sizeof(c));
    const int64_t N = cfg.n_agents, C = cfg.n_cells, M = cfg.max_intervals;
    c.n_agents = N; c.n_cells = C; c.cap_intervals = M;
    // rows: sum over cells of (k*m + m/2)/2 + 2 with k <= 19 -> generous static bound, clipped by config
    c.cap_scratch_rows = 10 * M + 2 * std::min(M, C) + 1024;
    c.cap_rows = cfg.max_rows > 0 ? cfg.max_rows : (4 * M + 1024);
    c.cap_expo = cfg.max_exposures > 0 ? cfg.max_exposures : std::max<int64_t>(N / 4, 4096);
    const int64_t S = std::min(M, C);   // max active cells
    int8_t *ct; int16_t *ci; int8_t *ab, *eb; double *sr; int32_t *rc;
    DA(ct, C); DA(ci, C); DA(ab, N); DA(eb, N); DA(sr, N); DA(rc, N);
    c.cell_type = ct; c.cell_indid = ci; c.age_bracket = ab; c.epi_age_bracket = eb; c.soc_rate = sr; c.res_cellid = rc;
    { AgentAttr *at; DA(at, N); c.k1_attr = at; }
    DA(c.seir, N); DA(c.inf_type, N); DA(c.inf_sev, N); DA(c.last_dose_day, N);
    DA(c.sched, N * MTD_MAXSCHED); DA(c.sched_cnt, N); DA(c.sched_next, N + 8); DA(c.test_day, 2 * N); DA(c.test_result_day, 2 * N);
    DA(c.quarantine, 3 * N); DA(c.hosp, 3 * N);
    DA(c.cell_count, C); DA(c.cell_first, C); DA(c.cell_base, C); DA(c.iv_pos, M); DA(c.flag_scan, M + 64);
    DA(c.slot_cell, S + 1); DA(c.slot_m, S + 1); DA(c.slot_first, S + 1); DA(c.slot_act, S + 1);
    DA(c.slot_off, S + 2); DA(c.slot_rowoff, S + 2); DA(c.slot_rowbound, S + 1);
    DA(c.slot_nrows, S + 1); DA(c.slot_denseoff, S + 2);
    c.class_stride = S + 1;
    DA(c.class_list, (MTD_NUM_CLASSES + 7) * (S + 1));      // + five regions: every class but the last again, sorted by size
    DA(c.k1_bins, 2 * 2 * 3 * 32); h->k1_bins_base = c.k1_bins;       // two sets: the small-cell ordering of two days may be in flight
    DA(c.csr, M);
    DA(c.scratch_rows, c.cap_scratch_rows); DA(c.rows, c.cap_rows);
    // cells above 1024 intervals are processed in a global workspace: as large as a cell of this handle can get
    c.max_cell_m = (int32_t)std::min<int64_t>(MTD_MAX_CELL_INTERVALS, M);
    h->k1_mm_big = c.max_cell_m > 1024 ? ((c.max_cell_m + 31) & ~31) : 0;
    if (h->k1_mm_big) DA(h->k1_big_ws, k1_big_ws_bytes(h->k1_mm_big));
    if (k1_use_legacy()) { c.k1_bits_words = k1_legacy_bits_words(h->sm_count); DA(c.k1_bits, c.k1_bits_words); }
    if (k1_use_split()) {
        for (int k = 0; k < 4; k++) {
            // the class streams carry the latency-bound chain kernels: above the throughput-bound small-cell kernels
            // (default priority), below the main / state streams
            int plo = 0, phi = 0;
            CK(cudaDeviceGetStreamPriorityRange(&plo, &phi));
            const char *pe = getenv("MTD_K1_CHAIN_PRIO");
            const int want = (pe && *pe) ? atoi(pe) : -1;
            const int cprio = want < phi ? phi : (want > plo ? plo : want);
            CK(cudaStreamCreateWithPriority(&h->k1cs.st[k], cudaStreamNonBlocking, cprio));
            const char *e3 = getenv("MTD_K1_CHAIN_PARITY");
            if (!(e3 && *e3 == '0')) {
                CK(cudaStreamCreateWithPriority(&h->k1cs.st2[k], cudaStreamNonBlocking, cprio));
                CK(cudaStreamCreateWithFlags(&h->k1cs.pst[k], cudaStreamNonBlocking));
            }
            CK(cudaEventCreateWithFlags(&h->k1cs.prep[k], cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&h->k1cs.done[k], cudaEventDisableTiming));
        }
        h->have_k1cs = true;
    }
    { const char *e = getenv("MTD_PARITIES"); h->n_par = (e && *e == '2') ? 2 : 3; }
    if (k1_use_split()) { c.k1_split = 1; for (int q = 0; q < h->n_par; q++) DA(h->k1_rec[q], MTD_K1_REC_STRIDE * (size_t)M + 64); c.k1_rec = h->k1_rec[0]; }
    DA(c.win, N); DA(c.succ, c.cap_rows);
    c.row_p = nullptr;
    DA(h->delta_own, (c.cap_expo + 1) * MTD_DELTA_WORDS);
    c.delta = h->delta_own;
    DA(c.ctr, 1);
    DA(c.sctr, 1);
    DA(h->d_params, 1);
    c.P = h->d_params;
    {   // math.log(i) for small integers, from the host libm (what CPython's math.log calls)
        double *lt; DA(lt, MTD_LOGTAB_N);
        std::vector<double> tab(MTD_LOGTAB_N);
        tab[0] = -INFINITY;
        for (int i = 1; i < MTD_LOGTAB_N; i++) tab[i] = std::log((double)i);
        CK(cudaMemcpy(lt, tab.data(), sizeof(double) * MTD_LOGTAB_N, cudaMemcpyHostToDevice));
        c.logtab = lt;
    }
    DA(h->d_iv_agent, M); DA(h->d_iv_cell, M); DA(h->d_iv_start, M); DA(h->d_iv_end, M); DA(h->d_iv_order, M);
    DA(h->d_mask, N);
    DA(h->scan_tmp, 2 * ((std::max(M, S) + 2) / 2048 + 16));
    DA(h->scan_tmp_k0, 2 * ((std::max(M, S) + 2) / 2048 + 16));
    {   // the further sets of the K0 products (a day's set is reused n_par days later)
        mtd_handle_s::K0Set &a = h->k0set[0];
        a = { c.cell_count, c.cell_first, c.cell_base, c.iv_pos, c.flag_scan, c.slot_cell, c.slot_m, c.slot_first, c.slot_act,
              c.slot_off, c.slot_rowoff, c.slot_rowbound, c.slot_nrows, c.slot_denseoff, c.class_list, c.csr, c.ctr, c.scratch_rows };
        for (int q = 1; q < h->n_par; q++) {
            mtd_handle_s::K0Set &b = h->k0set[q];
            DA(b.cell_count, C); DA(b.cell_first, C); DA(b.cell_base, C); DA(b.iv_pos, M); DA(b.flag_scan, M + 64);
            DA(b.slot_cell, S + 1); DA(b.slot_m, S + 1); DA(b.slot_first, S + 1); DA(b.slot_act, S + 1);
            DA(b.slot_off, S + 2); DA(b.slot_rowoff, S + 2); DA(b.slot_rowbound, S + 1);
            DA(b.slot_nrows, S + 1); DA(b.slot_denseoff, S + 2);
            DA(b.class_list, (MTD_NUM_CLASSES + 7) * (S + 1));
            DA(b.csr, M); DA(b.ctr, 1); DA(b.scratch_rows, c.cap_scratch_rows);
            CK(cudaMemsetAsync(b.ctr, 0, sizeof(DevCounters), h->stream));
        }
        int lo = 0, hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        const char *pe = getenv("MTD_K0_PRIO");
        // K0 of day d+1 gates K1 of day d+1, and K1 stages hold the machine back to back: at the default priority K0 only got
        // its SM slots when K1 of day d drained, which put K1(d) -> K0(d+1) -> K1(d+1) in a row (8-rank shard: 0.755 -> 0.721
        // ms/day with K0 above the K1 kernels, profiles/r02_k1_schedule_ab.txt).  MTD_K0_PRIO=0: lowest priority.
        const int prio = (pe && *pe == '0') ? lo : hi;
        CK(cudaStreamCreateWithPriority(&h->k0_stream, cudaStreamNonBlocking, prio));
        CK(cudaEventCreateWithFlags(&h->ev_k0_ready, cudaEventDisableTiming));
        CK(cudaStreamCreateWithFlags(&h->k1_stream, cudaStreamNonBlocking));
        { const char *e4 = getenv("MTD_K1_STREAMS"); if (!(e4 && *e4 == '1')) CK(cudaStreamCreateWithFlags(&h->k1_stream2, cudaStreamNonBlocking)); }
        CK(cudaEventCreateWithFlags(&h->ev_k1_done, cudaEventDisableTiming));
        { const char *e1 = getenv("MTD_K1_PIPE"); h->k1_pipe = !(e1 && *e1 == '0'); }
        if (h->have_k1cs) {
            CK(cudaStreamCreateWithFlags(&h->k1_join_stream, cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&h->ev_k1_small, cudaEventDisableTiming));
            const char *e2 = getenv("MTD_K1_DAYS_OVERLAP"); h->k1_days_overlap = !(e2 && *e2 == '0');
        }
        for (int q = 0; q < MTD_MAX_PARITIES; q++) CK(cudaEventCreateWithFlags(&h->ev_main_mark[q], cudaEventDisableTiming));
        DA(h->n_succ_buf, MTD_MAX_PARITIES);
        CK(cudaMemsetAsync(h->n_succ_buf, 0, MTD_MAX_PARITIES * sizeof(int32_t), h->stream));
        c.n_succ_pub = h->n_succ_buf;
        const char *e = getenv("MTD_K0_PIPE");
        h->k0_pipe = !(e && *e == '0');
        const char *es = getenv("MTD_K0_PIPE_STAGED");
        h->k0_pipe_staged = es && *es == '1';
    }
    DA(h->expo_list, 2 * c.cap_expo);
    DA(h->occ, 8);
    // initial state: nobody present
    CK(cudaMemsetAsync(c.seir, 0, N, h->stream)); CK(cudaMemsetAsync(c.inf_type, 0, N, h->stream));
    CK(cudaMemsetAsync(c.inf_sev, 0, N, h->stream)); CK(cudaMemsetAsync(c.sched_cnt, 0, N, h->stream));
    CK(cudaMemsetAsync(c.sched, 0, sizeof(int32_t) * N * MTD_MAXSCHED, h->stream));
    CK(cudaMemsetAsync(c.sched_next, 0x7f, sizeof(int16_t) * (N + 8), h->stream));      // MTD_SCHED_NEXT_NONE
    CK(cudaMemsetAsync(c.ctr, 0, sizeof(DevCounters), h->stream));
    CK(cudaMemsetAsync(c.sctr, 0, sizeof(StateCounters), h->stream));
    CK(cudaMemsetAsync(ci, 0xff, sizeof(int16_t) * C, h->stream));
    std::vector<int32_t> none((size_t)std::max<int64_t>(3 * N, 1), MTD_NONE);
    CK(cudaMemcpyAsync(c.last_dose_day, none.data(), sizeof(int32_t) * N, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(c.test_day, none.data(), sizeof(int32_t) * 2 * N, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(c.test_result_day, none.data(), sizeof(int32_t) * 2 * N, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(c.quarantine, none.data(), sizeof(int32_t) * 3 * N, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(c.hosp, none.data(), sizeof(int32_t) * 3 * N, cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return MTD_OK;
}

extern "C" int mtd_create(const mtd_config *cfg, mtd_handle *out)
{
    if (!cfg || !out || cfg->n_agents <= 0 || cfg->n_cells <= 0 || cfg->max_intervals <= 0) return MTD_E_ARG;
    mtd_handle h = new (std::nothrow) mtd_handle_s();
    if (!h) return MTD_E_ARG;
    h->cfg = *cfg;
    *out = h;                // returned even on failure so that mtd_last_error() can be read; caller destroys it
    return create_impl(h);
}

static void free_tape(mtd_handle h)
{
    for (void *p : h->tape_allocs) cudaFree(p);
    h->tape_allocs.clear();
    memset(&h->c.tape, 0, sizeof(h->c.tape));
}

extern "C" int mtd_destroy(mtd_handle h)
{
    if (!h) return MTD_E_ARG;
    cudaSetDevice(h->cfg.device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->state_stream) { cudaStreamSynchronize(h->state_stream); cudaStreamDestroy(h->state_stream); }
    if (h->k0_stream) { cudaStreamSynchronize(h->k0_stream); cudaStreamDestroy(h->k0_stream); }
    if (h->k1_stream) { cudaStreamSynchronize(h->k1_stream); cudaStreamDestroy(h->k1_stream); }
    if (h->k1_stream2) { cudaStreamSynchronize(h->k1_stream2); cudaStreamDestroy(h->k1_stream2); }
    if (h->k1_join_stream) { cudaStreamSynchronize(h->k1_join_stream); cudaStreamDestroy(h->k1_join_stream); }
    if (h->ev_k1_small) cudaEventDestroy(h->ev_k1_small);
    if (h->ev_k1_done) cudaEventDestroy(h->ev_k1_done);
    for (cudaEvent_t e : { h->ev_k0_ready, h->ev_main_mark[0], h->ev_main_mark[1], h->ev_main_mark[2] }) if (e) cudaEventDestroy(e);
    if (h->ev_state) cudaEventDestroy(h->ev_state);
    if (h->ev_main) cudaEventDestroy(h->ev_main);
    for (void *p : { (void *)h->k6_agent, (void *)h->k6_u }) if (p) cudaFree(p);
    delete h->trace_ctx;
    for (void *p : { (void *)h->seed_buf, (void *)h->sc_a1, (void *)h->sc_a2, (void *)h->sc_off, (void *)h->sc_se, (void *)h->vacc_ids, (void *)h->vacc_ts,
                     (void *)h->k8_ent_off, (void *)h->k8_grp_off, (void *)h->k8_out_off, (void *)h->k8_ent_ts, (void *)h->k8_ent_cell,
                     (void *)h->k8_grp_agent, (void *)h->k8_grp_res, (void *)h->k8_count, (void *)h->k8_scan }) if (p) cudaFree(p);
    for (void *p : h->xchg_peers) cudaIpcCloseMemHandle(p);
    if (h->xchg_alloc) cudaFree(h->xchg_alloc);
    free_tape(h);
    for (void *p : h->trace_tape_allocs) cudaFree(p);
    for (void *p : { (void *)h->res_off, (void *)h->res_nres, (void *)h->res_members, (void *)h->trace_sel, (void *)h->res_first,
                     (void *)h->res_k, (void *)h->is_traced, (void *)h->primary, (void *)h->secondary, (void *)h->d_epi_present,
                     (void *)h->trace_order, (void *)h->trace_cursor, (void *)h->sec_owner, (void *)h->trace_scalars,
                     (void *)h->trace_flagscan, (void *)h->trace_scan_tmp, (void *)h->d_traced, (void *)h->d_rows3 })
        if (p) cudaFree(p);
    for (int s = 0; s < MTD_N_STAGES; s++) for (cudaEvent_t e : h->ev[s]) cudaEventDestroy(e);
    if (h->have_k1cs) for (int k = 0; k < 4; k++) { cudaStreamSynchronize(h->k1cs.st[k]); cudaStreamDestroy(h->k1cs.st[k]);
        if (h->k1cs.st2[k]) { cudaStreamSynchronize(h->k1cs.st2[k]); cudaStreamDestroy(h->k1cs.st2[k]); }
        if (h->k1cs.pst[k]) { cudaStreamSynchronize(h->k1cs.pst[k]); cudaStreamDestroy(h->k1cs.pst[k]); }
        cudaEventDestroy(h->k1cs.prep[k]); cudaEventDestroy(h->k1cs.done[k]); }
    if (h->aux_stream) { cudaStreamSynchronize(h->aux_stream); cudaStreamDestroy(h->aux_stream); }
    if (h->aux2_stream) { cudaStreamSynchronize(h->aux2_stream); cudaStreamDestroy(h->aux2_stream); }
    if (h->ev_join2) cudaEventDestroy(h->ev_join2);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    if (h->copy_stream) { cudaStreamSynchronize(h->copy_stream); cudaStreamDestroy(h->copy_stream); }
    if (h->export_stream) { cudaStreamSynchronize(h->export_stream); cudaStreamDestroy(h->export_stream); }
    if (h->bulk_stream) { cudaStreamSynchronize(h->bulk_stream); cudaStreamDestroy(h->bulk_stream); }
    for (cudaEvent_t e : { h->ev_staged[0], h->ev_staged[1], h->ev_k0_done, h->ev_snap, h->ev_exported[0], h->ev_exported[1],
                           h->ev_res[0], h->ev_res[1] })
        if (e) cudaEventDestroy(e);
    for (void *p : h->allocs) cudaFree(p);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return MTD_OK;
}

static cudaMemcpyKind kind_in(int mem) { return mem == MTD_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice; }
static cudaMemcpyKind kind_out(int mem) { return mem == MTD_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost; }

#define COPY_IN(dst, src, bytes) do { if (src) CK(cudaMemcpyAsync((void *)(dst), (src), (bytes), kind_in(mem), h->stream)); } while (0)
#define COPY_OUT(dst, src, bytes) do { if (dst) CK(cudaMemcpyAsync((dst), (const void *)(src), (bytes), kind_out(mem), h->stream)); } while (0)

extern "C" int mtd_set_params(mtd_handle h, const mtd_params *p)
{
    if (!h || !p) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaMemcpyAsync(h->d_params, p, sizeof(mtd_params), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    h->have_params = true;
    return MTD_OK;
}

extern "C" int mtd_set_cells(mtd_handle h, int64_t n, const int8_t *cell_type, const int16_t *cell_indid, int mem)
{
    if (!h || n != h->c.n_cells || !cell_type) { if (h) h->err = "mtd_set_cells: bad arguments"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    COPY_IN(h->c.cell_type, cell_type, (size_t)n);
    COPY_IN(h->c.cell_indid, cell_indid, sizeof(int16_t) * (size_t)n);
    if (mem == MTD_MEM_HOST) CK(cudaStreamSynchronize(h->stream));
    h->have_cells = true;
    return MTD_OK;
}

extern "C" int mtd_set_static(mtd_handle h, int64_t n, const int8_t *age_bracket, const int8_t *epi_age_bracket,
                              const double *soc_rate, const int32_t *res_cellid, int mem)
{
    if (!h || n != h->c.n_agents) { if (h) h->err = "mtd_set_static: n != n_agents"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    COPY_IN(h->c.age_bracket, age_bracket, (size_t)n);
    COPY_IN(h->c.epi_age_bracket, epi_age_bracket, (size_t)n);
    COPY_IN(h->c.soc_rate, soc_rate, sizeof(double) * (size_t)n);
    COPY_IN(h->c.res_cellid, res_cellid, sizeof(int32_t) * (size_t)n);
    launch_pack_attr(h->c, const_cast<AgentAttr *>(h->c.k1_attr), h->stream, h->sm_count);
    if (mem == MTD_MEM_HOST) CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    h->have_static = true;
    return MTD_OK;
}

extern "C" int mtd_set_state(mtd_handle h, int64_t n, const int8_t *seir, const int8_t *inf_type, const int8_t *inf_sev,
                             const int32_t *last_dose_day, const int32_t *sched, const uint8_t *sched_cnt,
                             const int32_t *test_day, const int32_t *test_result_day, const int32_t *quarantine,
                             const int32_t *hosp, int mem)
{
    if (!h || n != h->c.n_agents) { if (h) h->err = "mtd_set_state: n != n_agents"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    size_t N = (size_t)n;
    COPY_IN(h->c.seir, seir, N); COPY_IN(h->c.inf_type, inf_type, N); COPY_IN(h->c.inf_sev, inf_sev, N);
    COPY_IN(h->c.last_dose_day, last_dose_day, 4 * N);
    COPY_IN(h->c.sched, sched, 4 * N * MTD_MAXSCHED); COPY_IN(h->c.sched_cnt, sched_cnt, N);
    COPY_IN(h->c.test_day, test_day, 8 * N); COPY_IN(h->c.test_result_day, test_result_day, 8 * N);
    COPY_IN(h->c.quarantine, quarantine, 12 * N); COPY_IN(h->c.hosp, hosp, 12 * N);
    if (sched || sched_cnt) launch_k4_rebuild_next(h->c, h->stream, h->sm_count);      // K4's per-agent "earliest entry day"
    if (mem == MTD_MEM_HOST) CK(cudaStreamSynchronize(h->stream));
    return MTD_OK;
}

extern "C" int mtd_get_state(mtd_handle h, int64_t n, int8_t *seir, int8_t *inf_type, int8_t *inf_sev, int32_t *sched,
                             uint8_t *sched_cnt, int32_t *test_day, int32_t *test_result_day, int32_t *quarantine,
                             int32_t *hosp, int mem)
{
    if (!h || n != h->c.n_agents) { if (h) h->err = "mtd_get_state: n != n_agents"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    size_t N = (size_t)n;
    COPY_OUT(seir, h->c.seir, N); COPY_OUT(inf_type, h->c.inf_type, N); COPY_OUT(inf_sev, h->c.inf_sev, N);
    COPY_OUT(sched, h->c.sched, 4 * N * MTD_MAXSCHED); COPY_OUT(sched_cnt, h->c.sched_cnt, N);
    COPY_OUT(test_day, h->c.test_day, 8 * N); COPY_OUT(test_result_day, h->c.test_result_day, 8 * N);
    COPY_OUT(quarantine, h->c.quarantine, 12 * N); COPY_OUT(hosp, h->c.hosp, 12 * N);
    CK(cudaStreamSynchronize(h->stream));
    return MTD_OK;
}

template <typename T>
static int tape_up(mtd_handle h, const T **dst, const T *src, size_t n)
{
    void *q = nullptr;
    cudaError_t e = cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T));
    if (e != cudaSuccess) { h->err = std::string("cudaMalloc(tape): ") + cudaGetErrorString(e); return MTD_E_CUDA; }
    h->tape_allocs.push_back(q);
    if (n) {
        e = cudaMemcpyAsync(q, src, n * sizeof(T), cudaMemcpyHostToDevice, h->stream);
        if (e != cudaSuccess) { h->err = std::string("cudaMemcpy(tape): ") + cudaGetErrorString(e); return MTD_E_CUDA; }
    }
    *dst = (const T *)q;
    return 0;
}

extern "C" int mtd_set_tape(mtd_handle h, const mtd_tape *t)
{
    if (!h) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    free_tape(h);
    if (!t) return MTD_OK;
    DevTape &d = h->c.tape;
    d.n_sel = t->n_sel; d.n_pair = t->n_pair; d.n_expo = t->n_expo;
    int r;
    if ((r = tape_up(h, &d.sel_key, t->sel_key, (size_t)t->n_sel))) return r;
    if ((r = tape_up(h, &d.sel_off, t->sel_off, (size_t)t->n_sel + 1))) return r;
    size_t npick = t->n_sel ? (size_t)t->sel_off[t->n_sel] : 0;
    if ((r = tape_up(h, &d.sel_picks, t->sel_picks, npick))) return r;
    if ((r = tape_up(h, &d.pair_k1, t->pair_k1, (size_t)t->n_pair))) return r;
    if ((r = tape_up(h, &d.pair_k2, t->pair_k2, (size_t)t->n_pair))) return r;
    if ((r = tape_up(h, &d.pair_u, t->pair_u, (size_t)t->n_pair))) return r;
    if ((r = tape_up(h, &d.expo_agent, t->expo_agent, (size_t)t->n_expo))) return r;
    if ((r = tape_up(h, &d.expo_row, t->expo_row, (size_t)t->n_expo * MTD_EXPO_W))) return r;
    CK(cudaStreamSynchronize(h->stream));
    return MTD_OK;
}

static int issue_host_export(mtd_handle h, int sl);

extern "C" int mtd_simulate_contact_network(mtd_handle h, const mtd_day *day, int64_t M, const int32_t *iv_agent,
                                            const int32_t *iv_cell, const uint8_t *iv_start, const uint8_t *iv_end,
                                            const uint32_t *iv_order, int mem)
{
    if (!h || !day || M < 0) return MTD_E_ARG;
    if (!(h->have_params && h->have_cells && h->have_static)) { h->err = "params / cells / static not set"; return MTD_E_STATE; }
    if (M > h->c.cap_intervals) { h->err = "interval stream exceeds max_intervals"; return MTD_E_CAPACITY; }
    if (day->mode == MTD_MODE_REPLAY && M > 0 && h->c.tape.sel_key == nullptr) { h->err = "replay mode without a tape"; return MTD_E_STATE; }
    CK(cudaSetDevice(h->cfg.device));
    DevCtx &c = h->c;
    c.day = *day;
    c.M = M;
    c.ov_off = nullptr; c.ov_se = nullptr;
    c.agent_masked = h->mask_on ? h->agent_masked : nullptr;
    if (h->mask_on) TOUCH();                 // K0 reads today's state (after K4 of today and the apply of yesterday)
    // today's parity of the K0 products; K0 goes to its own stream when its inputs do not come through the main stream
    const int kp = (int)(h->k0_seq++ % h->n_par);
    {
        const mtd_handle_s::K0Set &k = h->k0set[kp];
        c.cell_count = k.cell_count; c.cell_first = k.cell_first; c.cell_base = k.cell_base; c.iv_pos = k.iv_pos;
        c.flag_scan = k.flag_scan; c.slot_cell = k.slot_cell; c.slot_m = k.slot_m; c.slot_first = k.slot_first;
        c.slot_act = k.slot_act; c.slot_off = k.slot_off; c.slot_rowoff = k.slot_rowoff; c.slot_rowbound = k.slot_rowbound;
        c.slot_nrows = k.slot_nrows; c.slot_denseoff = k.slot_denseoff; c.class_list = k.class_list; c.csr = k.csr; c.ctr = k.ctr;
        c.scratch_rows = k.scratch_rows;
        c.n_succ_pub = h->n_succ_buf + kp;
        if (c.k1_split) c.k1_rec = h->k1_rec[kp];
    }
    const bool pipe = h->k0_pipe && !h->mask_on && M > 0 && ((mem == MTD_MEM_STAGED && h->k0_pipe_staged) || (mem == MTD_MEM_DEVICE && h->k0_device_ready));
    cudaStream_t ks = pipe ? h->k0_stream : h->stream;
    // this parity was last read by the day before yesterday's K1..K3 (main stream): all of it had been enqueued when
    // yesterday's call recorded its mark.  (The apply chain on the state stream reads no K0 product: the one counter it
    // needs, n_succ, is published by K3 into a slot of its own.)
    if (pipe) CK(cudaStreamWaitEvent(ks, h->ev_main_mark[(kp + 1) % h->n_par], 0));
    CK(cudaEventRecord(h->ev_main_mark[kp], h->stream));
    // first pipelined day after a serial one: that day's K0/K1 ran on the main stream and own the K1 scratch until they end
    if (pipe && !h->last_pipe) CK(cudaStreamWaitEvent(ks, h->ev_main_mark[kp], 0));
    h->last_pipe = pipe;
    if (mem == MTD_MEM_STAGED) {
        const int sl = h->st_ready;
        if (sl < 0 || h->st_n[sl] != M) { h->err = "MTD_MEM_STAGED without a matching mtd_prefetch_intervals"; return MTD_E_STATE; }
        CK(cudaStreamWaitEvent(ks, h->ev_staged[sl], 0));
        c.iv_agent = h->st_agent[sl]; c.iv_cell = h->st_cell[sl]; c.iv_start = h->st_start[sl]; c.iv_end = h->st_end[sl];
        c.iv_order = h->st_has_order[sl] ? h->st_order[sl] : nullptr;
        h->st_ready = -1;
    } else if (mem == MTD_MEM_BUILT) {
        if (h->built_n != M) { h->err = "MTD_MEM_BUILT without a matching mtd_build_intervals"; return MTD_E_STATE; }
        c.iv_agent = h->d_iv_agent; c.iv_cell = h->d_iv_cell; c.iv_start = h->d_iv_start; c.iv_end = h->d_iv_end; c.iv_order = nullptr;
    } else if (mem == MTD_MEM_HOST) {
        h->built_n = -1;                      // the staging buffers are overwritten
        StageScope sc(h, MTD_STAGE_H2D);
        size_t m = (size_t)M;
        CK(cudaMemcpyAsync(h->d_iv_agent, iv_agent, 4 * m, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->d_iv_cell, iv_cell, 4 * m, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->d_iv_start, iv_start, m, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->d_iv_end, iv_end, m, cudaMemcpyHostToDevice, h->stream));
        if (iv_order) CK(cudaMemcpyAsync(h->d_iv_order, iv_order, 4 * m, cudaMemcpyHostToDevice, h->stream));
        c.iv_agent = h->d_iv_agent; c.iv_cell = h->d_iv_cell; c.iv_start = h->d_iv_start; c.iv_end = h->d_iv_end;
        c.iv_order = iv_order ? h->d_iv_order : nullptr;
    } else {
        c.iv_agent = iv_agent; c.iv_cell = iv_cell; c.iv_start = iv_start; c.iv_end = iv_end; c.iv_order = iv_order;
    }
    CK(cudaMemsetAsync(c.ctr, 0, sizeof(DevCounters), ks));
    CK(cudaMemsetAsync(c.cell_count, 0, sizeof(int32_t) * (size_t)c.n_cells, ks));
    CK(cudaMemsetAsync(c.cell_first, 0xff, sizeof(uint32_t) * (size_t)c.n_cells, ks));
    { StageScope sc(h, MTD_STAGE_K0, ks); sc.add(launch_k0(c, pipe ? h->scan_tmp_k0 : h->scan_tmp, ks, h->sm_count)); }
    if (h->ev_k0_done) CK(cudaEventRecord(h->ev_k0_done, ks));             // the interval stream has been consumed
    // K1 needs K0 of its own day only: in the pipelined mode it runs on a third stream, beside K2/K3 of the day before
    // (main stream), and the main stream picks the day up again at K2
    const bool k1_alt = h->k1_stream2 && h->k1_join_stream && h->k1_days_overlap && ((h->k0_seq - 1) & 1);
    cudaStream_t k1s = (pipe && h->k1_pipe) ? (k1_alt ? h->k1_stream2 : h->k1_stream) : h->stream;
    c.k1_bins = h->k1_bins_base + (((h->k0_seq - 1) & 1) ? 2 * 3 * 32 : 0);
    if (pipe) { CK(cudaEventRecord(h->ev_k0_ready, ks)); CK(cudaStreamWaitEvent(k1s, h->ev_k0_ready, 0)); }
    // an empty day (M == 0) has no cells: the zeroed counters are its result
    // Day overlap of K1 (pipelined mode, class streams): the K1 stream does not wait for the stage's side streams -- the
    // chain kernels of the large classes are bound by the latency of their longest cell, and the next day's small-cell
    // and prepare kernels fill the machine beside them (every product of the stage is double-buffered by day parity, and
    // each class stream takes the days in order).  The stage is joined on a stream of its own, which the main stream waits
    // for before K2 and which carries the stage's end event.
    if (M > 0) {
        StageScope sc(h, MTD_STAGE_K1, k1s);
        bool joined = true;
        const bool overlap_days = k1s != h->stream && h->k1_join_stream && h->k1_days_overlap;
        h->k1cs.par = (int)((h->k0_seq - 1) & 1);      // chain streams alternate day by day, whatever the number of product sets
        sc.add(launch_k1(c, k1s, h->aux_stream, h->aux2_stream, h->ev_fork, h->ev_join, h->ev_join2, h->sm_count, h->k1_big_ws, h->k1_mm_big,
                         h->have_k1cs ? &h->k1cs : nullptr, overlap_days ? &joined : nullptr));
        if (!joined) {
            cudaStream_t js = h->k1_join_stream;
            CK(cudaEventRecord(h->ev_k1_small, k1s)); CK(cudaStreamWaitEvent(js, h->ev_k1_small, 0));
            CK(cudaStreamWaitEvent(js, h->ev_join, 0)); CK(cudaStreamWaitEvent(js, h->ev_join2, 0));
            for (int k = 0; k < 4; k++) CK(cudaStreamWaitEvent(js, h->k1cs.done[k], 0));
            sc.st = js;                                  // the stage ends when its last stream does
            CK(cudaEventRecord(h->ev_k1_done, js));
            CK(cudaStreamWaitEvent(h->stream, h->ev_k1_done, 0));
        } else if (k1s != h->stream) { CK(cudaEventRecord(h->ev_k1_done, k1s)); CK(cudaStreamWaitEvent(h->stream, h->ev_k1_done, 0)); }
    } else if (k1s != h->stream) { CK(cudaEventRecord(h->ev_k1_done, k1s)); CK(cudaStreamWaitEvent(h->stream, h->ev_k1_done, 0)); }
    c.rows12 = nullptr;
    if (h->export_on) {     // K2 also packs the rows for the export; the slot's previous export must have drained
        const int sl = (int)(h->day_seq & 1);
        { int r = issue_host_export(h, sl); if (r) return r; }
        if (h->exported_pending[sl]) { CK(cudaStreamWaitEvent(h->stream, h->ev_exported[sl], 0)); h->exported_pending[sl] = false; }
        c.rows12 = h->rows12_buf[sl];
    }
    h->day_seq++;
    // K2 overwrites the dense rows the apply chain of the day before still reads when it clears its winners
    { int r = main_after_state(h); if (r) return r; }
    if (M > 0) { StageScope sc(h, MTD_STAGE_K2); sc.add(launch_scan_nrows(c, h->scan_tmp, h->stream));
                 sc.add(launch_k2_compact(c, h->stream, h->sm_count)); }
    CK(cudaGetLastError());
    h->day_ready = true;
    return MTD_OK;
}

// apply chain of the day's exposure records on the state stream: [peer wait] -> [resolve] -> apply -> clear winners
static int enqueue_apply(mtd_handle h, const int32_t *bufs, int n_bufs, int64_t cap, bool resolve, bool wait_peers, int par)
{
    { int r = state_after_main(h); if (r) return r; }
    cudaStream_t ss = h->state_stream;
    CK(cudaMemsetAsync(h->c.sctr, 0, sizeof(StateCounters), ss));
    StageScope sc(h, MTD_STAGE_K3_APPLY, ss);
    if (wait_peers) sc.add(launch_xchg_wait(h->x, par, h->xchg_seq, h->c.sctr, ss));
    sc.add(launch_k3_apply(h->c, bufs, n_bufs, cap, resolve, h->expo_list, ss, h->sm_count));
    h->win_clean = true;
    h->state_dirty = true;
    return MTD_OK;
}

extern "C" int mtd_simulate_direct_contacts(mtd_handle h, const mtd_day *day, int32_t *delta_buf, int64_t cap, int apply)
{
    if (!h || !day) return MTD_E_ARG;
    if (!h->day_ready) { h->err = "mtd_simulate_direct_contacts before mtd_simulate_contact_network"; return MTD_E_STATE; }
    if (apply == MTD_APPLY_EXCHANGE && !h->xchg_on) { h->err = "MTD_APPLY_EXCHANGE before mtd_exchange_ready"; return MTD_E_STATE; }
    CK(cudaSetDevice(h->cfg.device));
    DevCtx &c = h->c;
    c.day = *day;
    int32_t *buf = delta_buf ? delta_buf : h->delta_own;
    int64_t cp = delta_buf ? cap : c.cap_expo;
    int par = 0;
    const int64_t slice = (h->x.cap + 1) * MTD_DELTA_WORDS;
    if (apply == MTD_APPLY_EXCHANGE) {          // records go straight into this rank's slice of its own exchange buffer
        par = (int)((h->xchg_seq + 1) & 1);
        buf = h->x.buf[h->x.rank] + ((int64_t)par * h->x.world + h->x.rank) * slice;
        cp = h->x.cap;
    }
    if (cp <= 0) { h->err = "delta buffer capacity must be positive"; return MTD_E_ARG; }
    { int r = main_after_state(h); if (r) return r; }          // K4 of today and the apply of yesterday
    { StageScope sc(h, MTD_STAGE_K3_EVAL); sc.add(launch_k3_eval(c, buf, cp, h->win_clean, h->stream, h->sm_count)); h->win_clean = false;
      if (apply == MTD_APPLY_EXCHANGE) { h->xchg_seq++; sc.add(launch_xchg_push(h->x, par, h->xchg_seq, h->stream)); } }
    if (apply == MTD_APPLY_EXCHANGE) {
        int r = enqueue_apply(h, h->x.buf[h->x.rank] + (int64_t)par * h->x.world * slice, h->x.world, h->x.cap, true, true, par);
        if (r) return r;
    } else if (apply) {
        int r = enqueue_apply(h, buf, 1, cp, false, false, 0);
        if (r) return r;
    } else h->main_dirty = true;
    CK(cudaGetLastError());
    return MTD_OK;
}

template <typename T>
static int grow_buf(mtd_handle h, T **p, int64_t *cap, int64_t need)
{
    if (need <= *cap) return 0;
    if (*p) cudaFree(*p);
    *p = nullptr; *cap = 0;
    void *q = nullptr;
    const int64_t c = need + need / 2 + 256;
    cudaError_t e = cudaMalloc(&q, (size_t)c * sizeof(T));
    if (e != cudaSuccess) { h->err = std::string("cudaMalloc: ") + cudaGetErrorString(e); return MTD_E_CUDA; }
    *p = (T *)q; *cap = c;
    return 0;
}

extern "C" int mtd_set_contacts(mtd_handle h, const mtd_day *day, int32_t cellid, int64_t n_rows, const int32_t *a1,
                                const int32_t *a2, const int64_t *ov_off, const uint8_t *ov_se, int mem)
{
    if (!h || !day || n_rows < 0 || cellid < 0 || cellid >= h->c.n_cells || (n_rows && (!a1 || !a2 || !ov_off || !ov_se))) {
        if (h) h->err = "mtd_set_contacts: bad arguments"; return MTD_E_ARG; }
    if (mem != MTD_MEM_HOST) { h->err = "mtd_set_contacts: host buffers only"; return MTD_E_ARG; }
    if (!(h->have_params && h->have_cells && h->have_static)) { h->err = "params / cells / static not set"; return MTD_E_STATE; }
    if (n_rows > h->c.cap_rows) { h->err = "mtd_set_contacts: more rows than max_rows"; return MTD_E_CAPACITY; }
    CK(cudaSetDevice(h->cfg.device));
    DevCtx &c = h->c;
    c.day = *day;
    c.M = 0;
    const int64_t total = n_rows ? ov_off[n_rows] : 0;
    int64_t cap2 = h->sc_cap_rows;
    { int r; if ((r = grow_buf(h, &h->sc_a1, &h->sc_cap_rows, n_rows + 1))) return r;
      if ((r = grow_buf(h, &h->sc_a2, &cap2, n_rows + 1))) return r;
      if ((r = grow_buf(h, &h->sc_off, &h->sc_cap_off, n_rows + 2))) return r;
      if ((r = grow_buf(h, &h->sc_se, &h->sc_cap_se, 2 * total + 2))) return r; }
    CK(cudaMemsetAsync(c.ctr, 0, sizeof(DevCounters), h->stream));
    if (n_rows) {
        CK(cudaMemcpyAsync(h->sc_a1, a1, 4 * (size_t)n_rows, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->sc_a2, a2, 4 * (size_t)n_rows, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->sc_off, ov_off, 8 * (size_t)(n_rows + 1), cudaMemcpyHostToDevice, h->stream));
        if (total) CK(cudaMemcpyAsync(h->sc_se, ov_se, 2 * (size_t)total, cudaMemcpyHostToDevice, h->stream));
    }
    c.ov_off = h->sc_off; c.ov_se = h->sc_se;
    launch_set_contacts(c, cellid, n_rows, h->sc_a1, h->sc_a2, h->stream);
    CK(cudaStreamSynchronize(h->stream));          // the host buffers may be released by the caller
    CK(cudaGetLastError());
    h->day_ready = true;
    return MTD_OK;
}

extern "C" int mtd_seed_exposures(mtd_handle h, const mtd_day *day, int64_t n, const int32_t *agent_ids, const int32_t *exposure_day,
                                  int32_t pre_exposure_days, const uint8_t *timesteps, int32_t n_timesteps, const double *asym_mult,
                                  const double *draws, int32_t *out8)
{
    if (!h || !day || n < 0 || (n && (!agent_ids || !timesteps)) || n_timesteps <= 0) { if (h) h->err = "mtd_seed_exposures: bad arguments"; return MTD_E_ARG; }
    if (day->mode == MTD_MODE_REPLAY && n && (!draws || !exposure_day)) { h->err = "mtd_seed_exposures: replay mode needs draws and exposure days"; return MTD_E_ARG; }
    if (!exposure_day && pre_exposure_days <= 0) { h->err = "mtd_seed_exposures: pre_exposure_days must be positive"; return MTD_E_ARG; }
    if (!(h->have_params && h->have_static)) { h->err = "params / static not set"; return MTD_E_STATE; }
    if (n == 0) return MTD_OK;
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    h->c.day = *day;
    const size_t N = (size_t)n, W = MTD_EXPO_W;
    const size_t o_draws = 0, o_asym = o_draws + 8 * W * N, o_out = o_asym + 8 * N, o_ids = o_out + 32 * N, o_day = o_ids + 4 * N,
                 o_status = o_day + 4 * N, o_ts = o_status + 16, total = o_ts + (size_t)n_timesteps + 16;
    { int r = grow_buf(h, &h->seed_buf, &h->seed_cap, (int64_t)total); if (r) return r; }
    uint8_t *B = h->seed_buf;
    CK(cudaMemsetAsync(B + o_status, 0, 16, h->stream));
    CK(cudaMemcpyAsync(B + o_ids, agent_ids, 4 * N, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(B + o_ts, timesteps, (size_t)n_timesteps, cudaMemcpyHostToDevice, h->stream));
    if (exposure_day) CK(cudaMemcpyAsync(B + o_day, exposure_day, 4 * N, cudaMemcpyHostToDevice, h->stream));
    if (asym_mult) CK(cudaMemcpyAsync(B + o_asym, asym_mult, 8 * N, cudaMemcpyHostToDevice, h->stream));
    if (draws) CK(cudaMemcpyAsync(B + o_draws, draws, 8 * W * N, cudaMemcpyHostToDevice, h->stream));
    { StageScope sc(h, MTD_STAGE_K3_EVAL);
      sc.add(launch_k9_seed(h->c, n, (const int32_t *)(B + o_ids), exposure_day ? (const int32_t *)(B + o_day) : nullptr, pre_exposure_days,
                            B + o_ts, n_timesteps, asym_mult ? (const double *)(B + o_asym) : nullptr,
                            draws ? (const double *)(B + o_draws) : nullptr, (int32_t *)(B + o_out), (int32_t *)(B + o_status), h->stream)); }
    int32_t status[4] = { 0, 0, 0, 0 };
    CK(cudaMemcpyAsync(status, B + o_status, 16, cudaMemcpyDeviceToHost, h->stream));
    if (out8) CK(cudaMemcpyAsync(out8, B + o_out, 32 * N, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    if (status[0]) { h->err = "mtd_seed_exposures: agent id out of range"; return status[0]; }
    if (status[1]) { h->err = "mtd_seed_exposures: schedule overflow (MTD_MAXSCHED entries per agent)"; return MTD_E_CAPACITY; }
    if (status[2]) { h->err = "mtd_seed_exposures: replay draws missing (NaN slot read)"; return MTD_E_STATE; }
    return MTD_OK;
}

extern "C" int mtd_apply_deltas(mtd_handle h, const int32_t *bufs, int64_t n_bufs, int64_t cap)
{
    if (!h || !bufs || n_bufs <= 0 || cap <= 0) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    { int r = enqueue_apply(h, bufs, (int)n_bufs, cap, true, false, 0); if (r) return r; }
    CK(cudaGetLastError());
    return MTD_OK;
}

// ------------------------------------------------------------------------------------------------ K8 interval builder
extern "C" int mtd_build_intervals(mtd_handle h, int64_t n_itin, const int64_t *ent_off, const int16_t *ent_ts,
                                   const int32_t *ent_cell, const int64_t *grp_off, const int32_t *grp_agent,
                                   const int32_t *grp_rescell, int mem, int64_t *n_intervals)
{
    if (!h || n_itin < 0 || !n_intervals || (n_itin && (!ent_off || !ent_ts || !ent_cell || !grp_off || !grp_agent || !grp_rescell))) {
        if (h) h->err = "mtd_build_intervals: bad arguments"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    *n_intervals = 0; h->built_n = -1;
    if (n_itin == 0) { h->built_n = 0; return MTD_OK; }
    const int64_t *d_eo = ent_off, *d_go = grp_off; const int16_t *d_ts = ent_ts;
    const int32_t *d_ec = ent_cell, *d_ga = grp_agent, *d_gr = grp_rescell;
    int r;
    if (mem == MTD_MEM_HOST) {
        const int64_t n_ent = ent_off[n_itin], n_grp = grp_off[n_itin];
        if ((r = grow_buf(h, &h->k8_ent_off, &h->k8_cap[0], n_itin + 1))) return r;
        if ((r = grow_buf(h, &h->k8_grp_off, &h->k8_cap[1], n_itin + 1))) return r;
        if ((r = grow_buf(h, &h->k8_ent_ts, &h->k8_cap[2], n_ent + 1))) return r;
        if ((r = grow_buf(h, &h->k8_ent_cell, &h->k8_cap[3], n_ent + 1))) return r;
        if ((r = grow_buf(h, &h->k8_grp_agent, &h->k8_cap[4], n_grp + 1))) return r;
        if ((r = grow_buf(h, &h->k8_grp_res, &h->k8_cap[5], n_grp + 1))) return r;
        CK(cudaMemcpyAsync(h->k8_ent_off, ent_off, 8 * (size_t)(n_itin + 1), cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->k8_grp_off, grp_off, 8 * (size_t)(n_itin + 1), cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->k8_ent_ts, ent_ts, 2 * (size_t)n_ent, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->k8_ent_cell, ent_cell, 4 * (size_t)n_ent, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->k8_grp_agent, grp_agent, 4 * (size_t)n_grp, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(h->k8_grp_res, grp_rescell, 4 * (size_t)n_grp, cudaMemcpyHostToDevice, h->stream));
        d_eo = h->k8_ent_off; d_go = h->k8_grp_off; d_ts = h->k8_ent_ts; d_ec = h->k8_ent_cell; d_ga = h->k8_grp_agent; d_gr = h->k8_grp_res;
    }
    if ((r = grow_buf(h, &h->k8_count, &h->k8_cap[6], n_itin + 1))) return r;
    if ((r = grow_buf(h, &h->k8_out_off, &h->k8_cap[7], n_itin + 2))) return r;
    if ((r = grow_buf(h, &h->k8_scan, &h->k8_cap[8], 2 * ((n_itin + 2) / 2048 + 16)))) return r;
    int *d_err = (int *)(h->occ + 7);          // spare word behind the six occupancy counters
    CK(cudaMemsetAsync(d_err, 0, sizeof(int), h->stream));
    launch_k8(n_itin, d_eo, d_ts, d_ec, d_go, d_ga, d_gr, h->k8_count, h->k8_out_off, h->k8_scan, h->d_iv_agent, h->d_iv_cell,
              h->d_iv_start, h->d_iv_end, h->c.cap_intervals, d_err, h->stream, h->sm_count);
    int64_t total = 0; int err = 0;
    CK(cudaMemcpyAsync(&total, h->k8_out_off + n_itin, sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&err, d_err, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    *n_intervals = total;
    if (err || total > h->c.cap_intervals) { h->err = "mtd_build_intervals: the built stream exceeds max_intervals"; return MTD_E_CAPACITY; }
    h->built_n = total;
    return MTD_OK;
}

extern "C" int mtd_get_intervals(mtd_handle h, int64_t n, int32_t *iv_agent, int32_t *iv_cell, uint8_t *iv_start, uint8_t *iv_end, int mem)
{
    if (!h || n < 0 || n != h->built_n) { if (h) h->err = "mtd_get_intervals: n does not match the built stream"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    COPY_OUT(iv_agent, h->d_iv_agent, 4 * (size_t)n); COPY_OUT(iv_cell, h->d_iv_cell, 4 * (size_t)n);
    COPY_OUT(iv_start, h->d_iv_start, (size_t)n); COPY_OUT(iv_end, h->d_iv_end, (size_t)n);
    CK(cudaStreamSynchronize(h->stream));
    return MTD_OK;
}

// ------------------------------------------------------------------------------------------------ K7 vaccination
extern "C" int mtd_enable_vaccination(mtd_handle h, int64_t n_locals)
{
    if (!h || n_locals <= 0 || n_locals > h->c.n_agents) { if (h) h->err = "mtd_enable_vaccination: bad n_locals"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    if (!h->vacc_day) {
        DA(h->vacc_day, 2 * h->c.n_agents); DA(h->vacc_selected, h->c.n_agents);
        char *scr; DA(scr, 128); h->vacc_scratch = scr;
        DA(h->vacc_hist, 65536);
        TOUCH();
        std::vector<int32_t> none((size_t)(2 * h->c.n_agents), MTD_NONE);
        CK(cudaMemcpyAsync(h->vacc_day, none.data(), sizeof(int32_t) * none.size(), cudaMemcpyHostToDevice, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    h->vacc_n_locals = n_locals;
    return MTD_OK;
}

extern "C" int mtd_set_vaccination(mtd_handle h, int64_t n, const int32_t *last_dose_day, const int32_t *vacc_day, int mem)
{
    if (!h || n != h->c.n_agents) { if (h) h->err = "mtd_set_vaccination: n != n_agents"; return MTD_E_ARG; }
    if (!h->vacc_day) { h->err = "mtd_set_vaccination before mtd_enable_vaccination"; return MTD_E_STATE; }
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    COPY_IN(h->c.last_dose_day, last_dose_day, 4 * (size_t)n);
    COPY_IN(h->vacc_day, vacc_day, 8 * (size_t)n);
    if (mem == MTD_MEM_HOST) CK(cudaStreamSynchronize(h->stream));
    return MTD_OK;
}

extern "C" int mtd_get_vaccination(mtd_handle h, int64_t n, int32_t *last_dose_day, int32_t *vacc_day, int mem)
{
    if (!h || n != h->c.n_agents) { if (h) h->err = "mtd_get_vaccination: n != n_agents"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    COPY_OUT(last_dose_day, h->c.last_dose_day, 4 * (size_t)n);
    if (vacc_day) {
        if (!h->vacc_day) { h->err = "mtd_get_vaccination: vaccination is not enabled"; return MTD_E_STATE; }
        COPY_OUT(vacc_day, h->vacc_day, 8 * (size_t)n);
    }
    CK(cudaStreamSynchronize(h->stream));
    return MTD_OK;
}

extern "C" int mtd_schedule_vaccinations(mtd_handle h, const mtd_day *day, int64_t k, int64_t n_replay, const int32_t *replay_ids,
                                         const int32_t *replay_ts, uint8_t *selected, int64_t *n_selected, int mem)
{
    if (!h || !day || k < 0 || n_replay < 0) { if (h) h->err = "mtd_schedule_vaccinations: bad arguments"; return MTD_E_ARG; }
    if (!h->vacc_day) { h->err = "mtd_schedule_vaccinations before mtd_enable_vaccination"; return MTD_E_STATE; }
    if (day->mode == MTD_MODE_REPLAY && n_replay && (!replay_ids || !replay_ts)) { h->err = "replay mode without the reference's sample"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    h->c.day = *day;
    if (day->mode == MTD_MODE_REPLAY) {
        const int32_t *ids = replay_ids, *ts = replay_ts;
        if (mem == MTD_MEM_HOST && n_replay) {
            { int r; if ((r = grow_buf(h, &h->vacc_ids, &h->vacc_cap_ids, n_replay))) return r;
              if ((r = grow_buf(h, &h->vacc_ts, &h->vacc_cap_ts, n_replay))) return r; }
            CK(cudaMemcpyAsync(h->vacc_ids, replay_ids, 4 * (size_t)n_replay, cudaMemcpyHostToDevice, h->stream));
            CK(cudaMemcpyAsync(h->vacc_ts, replay_ts, 4 * (size_t)n_replay, cudaMemcpyHostToDevice, h->stream));
            ids = h->vacc_ids; ts = h->vacc_ts;
        }
        CK(cudaMemsetAsync(h->vacc_selected, 0, (size_t)h->c.n_agents, h->stream));
        launch_k7_replay(h->c, n_replay, ids, ts, h->vacc_day, h->vacc_selected, h->vacc_scratch, h->stream);
    } else {
        CK(cudaMemsetAsync(h->vacc_selected, 0, (size_t)h->c.n_agents, h->stream));
        launch_k7_schedule(h->c, h->vacc_n_locals, (long long)k, h->vacc_day, h->vacc_selected, h->vacc_scratch, h->vacc_hist,
                           h->stream, h->sm_count);
    }
    int cnt = 0;
    CK(cudaMemcpyAsync(&cnt, (char *)h->vacc_scratch + 64, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    if (selected) CK(cudaMemcpyAsync(selected, h->vacc_selected, (size_t)h->c.n_agents, kind_out(mem), h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (n_selected) *n_selected = cnt;
    CK(cudaGetLastError());
    return MTD_OK;
}

extern "C" int mtd_set_intervention_mask(mtd_handle h, int on)
{
    if (!h) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    if (on && !h->agent_masked) DA(h->agent_masked, h->c.n_agents);
    h->mask_on = on != 0;
    return MTD_OK;
}

extern "C" int mtd_set_k0_overlap(mtd_handle h, int on)
{
    if (!h) return MTD_E_ARG;
    h->k0_device_ready = on != 0;
    return MTD_OK;
}

extern "C" int mtd_advance_seir(mtd_handle h, int32_t day, const uint8_t *mask, int mem)
{
    if (!h) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    cudaStream_t ss = h->state_stream;
    if (h->main_dirty) { int r = state_after_main(h); if (r) return r; }
    const uint8_t *dm = mask;
    if (mask && mem == MTD_MEM_HOST) {
        CK(cudaMemcpyAsync(h->d_mask, mask, (size_t)h->c.n_agents, cudaMemcpyHostToDevice, ss));
        dm = h->d_mask;
    }
    { StageScope sc(h, MTD_STAGE_K4, ss); sc.add(launch_k4(h->c, day, dm, ss, h->sm_count));
      if (h->vacc_day) sc.add(launch_k7_dose(h->c, day, h->vacc_n_locals, h->vacc_day, ss, h->sm_count)); }
    h->state_dirty = true;
    CK(cudaGetLastError());
    return MTD_OK;
}

// ---- peer exchange: one allocation per rank = arrival flags [2][world] (256 bytes reserved) + slices [2][world][cap+1]
#define XCHG_FLAG_BYTES 256
extern "C" int mtd_exchange_init(mtd_handle h, int rank, int world, int64_t cap_records, void *ipc_handle_out)
{
    if (!h || rank < 0 || world < 2 || rank >= world || world > MTD_MAX_WORLD || cap_records <= 0 || !ipc_handle_out) {
        if (h) h->err = "mtd_exchange_init: bad arguments"; return MTD_E_ARG; }
    if (h->xchg_alloc) { h->err = "mtd_exchange_init: already initialised"; return MTD_E_STATE; }
    CK(cudaSetDevice(h->cfg.device));
    const size_t bytes = XCHG_FLAG_BYTES + sizeof(int32_t) * 2 * (size_t)world * (size_t)(cap_records + 1) * MTD_DELTA_WORDS;
    CK(cudaMalloc(&h->xchg_alloc, bytes));
    CK(cudaMemset(h->xchg_alloc, 0, bytes));
    CK(cudaDeviceSynchronize());
    h->x.rank = rank; h->x.world = world; h->x.cap = cap_records;
    h->x.flags[rank] = (unsigned long long *)h->xchg_alloc;
    h->x.buf[rank] = (int32_t *)((char *)h->xchg_alloc + XCHG_FLAG_BYTES);
    cudaIpcMemHandle_t ipc;
    CK(cudaIpcGetMemHandle(&ipc, h->xchg_alloc));
    memcpy(ipc_handle_out, &ipc, sizeof(ipc));
    return MTD_OK;
}

extern "C" int mtd_exchange_open(mtd_handle h, int peer, const void *ipc_handle)
{
    if (!h || !h->xchg_alloc || peer < 0 || peer >= h->x.world || peer == h->x.rank || !ipc_handle) {
        if (h) h->err = "mtd_exchange_open: bad arguments"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    cudaIpcMemHandle_t ipc;
    memcpy(&ipc, ipc_handle, sizeof(ipc));
    void *p = nullptr;
    CK(cudaIpcOpenMemHandle(&p, ipc, cudaIpcMemLazyEnablePeerAccess));
    h->xchg_peers.push_back(p);
    h->x.flags[peer] = (unsigned long long *)p;
    h->x.buf[peer] = (int32_t *)((char *)p + XCHG_FLAG_BYTES);
    return MTD_OK;
}

extern "C" int mtd_exchange_ready(mtd_handle h)
{
    if (!h || !h->xchg_alloc) { if (h) h->err = "mtd_exchange_ready before mtd_exchange_init"; return MTD_E_STATE; }
    for (int r = 0; r < h->x.world; r++)
        if (!h->x.buf[r]) { h->err = "mtd_exchange_ready: a peer's buffer has not been opened"; return MTD_E_STATE; }
    h->xchg_on = true;
    return MTD_OK;
}

extern "C" int mtd_set_profiling(mtd_handle h, int on)
{
    if (!h) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    if (on && h->ev[0].empty())
        for (int s = 0; s < MTD_N_STAGES; s++) {
            h->ev[s].resize(2 * mtd_handle_s::EV_RING);
            for (cudaEvent_t &e : h->ev[s]) CK(cudaEventCreate(&e));
        }
    h->profiling = on != 0;
    for (int s = 0; s < MTD_N_STAGES; s++) { h->ev_n[s] = 0; h->launches[s] = 0; }
    return MTD_OK;
}

extern "C" int mtd_get_stage_ms(mtd_handle h, float ms[MTD_N_STAGES], int64_t launches[MTD_N_STAGES])
{
    if (!h || !ms) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaStreamSynchronize(h->state_stream));
    CK(cudaStreamSynchronize(h->k0_stream));
    CK(cudaStreamSynchronize(h->k1_stream));
    if (h->k1_stream2) CK(cudaStreamSynchronize(h->k1_stream2));
    for (int s = 0; s < MTD_N_STAGES; s++) {
        ms[s] = 0.f;
        double sum = 0.0;
        for (int k = 0; k < h->ev_n[s]; k++) {
            float t = 0.f;
            CK(cudaEventElapsedTime(&t, h->ev[s][2 * k], h->ev[s][2 * k + 1]));
            sum += t;
        }
        if (h->ev_n[s] > 0) ms[s] = (float)(sum / h->ev_n[s]);
        h->ev_n[s] = 0;
        if (launches) launches[s] = h->launches[s];
        h->launches[s] = 0;
    }
    return MTD_OK;
}

extern "C" int mtd_sync(mtd_handle h)
{
    if (!h) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaStreamSynchronize(h->state_stream));
    CK(cudaStreamSynchronize(h->k0_stream));
    CK(cudaStreamSynchronize(h->k1_stream));
    if (h->k1_stream2) CK(cudaStreamSynchronize(h->k1_stream2));
    return MTD_OK;
}

extern "C" int mtd_get_counters(mtd_handle h, mtd_counters *out)
{
    if (!h || !out) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    DevCounters d; StateCounters sc;
    CK(cudaMemcpyAsync(&d, h->c.ctr, sizeof(d), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&sc, h->c.sctr, sizeof(sc), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    memset(out, 0, sizeof(*out));
    if (!d.error && sc.error) d.error = sc.error;
    out->n_slots = d.n_slots; out->n_rows = d.n_rows; out->n_exposed = sc.n_applied;
    out->pairs_evaluated = (int64_t)d.pairs_evaluated; out->n_potential = (int64_t)d.n_potential;
    out->n_incidences = (int64_t)d.n_incidences; out->n_si_rows = d.n_si_rows; out->sched_overflow = sc.sched_overflow; out->tape_miss = d.tape_miss;
    for (int k = 0; k < MTD_NUM_CLASSES; k++) out->class_cells[k] = d.class_cells[k];
    out->error = d.error;
    if (d.error) {
        char buf[256];
        snprintf(buf, sizeof(buf), "device error %d (rows needed %llu of %lld scratch / %d of %lld dense)",
                 d.error, d.row_bound_total, (long long)h->c.cap_scratch_rows, d.n_rows, (long long)h->c.cap_rows);
        h->err = buf;
        return d.error;
    }
    return MTD_OK;
}

extern "C" int mtd_get_contacts(mtd_handle h, int32_t *rows5, int32_t *row_dur, int32_t *row_cell, int64_t cap,
                                int64_t *n, int mem)
{
    if (!h || !n) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    DevCounters d;
    CK(cudaMemcpyAsync(&d, h->c.ctr, sizeof(d), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    *n = d.n_rows;
    if (d.n_rows > cap) { h->err = "mtd_get_contacts: cap too small"; return MTD_E_CAPACITY; }
    if (d.n_rows == 0) return MTD_OK;
    if (mem == MTD_MEM_DEVICE) {
        launch_export_rows(h->c, rows5, row_dur, row_cell, cap, h->stream);
        CK(cudaGetLastError());
        return MTD_OK;
    }
    int32_t *t5 = nullptr, *td = nullptr, *tc = nullptr;
    size_t R = (size_t)d.n_rows;
    CK(cudaMalloc((void **)&t5, R * 5 * 4)); CK(cudaMalloc((void **)&td, R * 4)); CK(cudaMalloc((void **)&tc, R * 4));
    launch_export_rows(h->c, t5, td, tc, (int64_t)R, h->stream);
    if (rows5) cudaMemcpyAsync(rows5, t5, R * 5 * 4, cudaMemcpyDeviceToHost, h->stream);
    if (row_dur) cudaMemcpyAsync(row_dur, td, R * 4, cudaMemcpyDeviceToHost, h->stream);
    if (row_cell) cudaMemcpyAsync(row_cell, tc, R * 4, cudaMemcpyDeviceToHost, h->stream);
    cudaError_t e = cudaStreamSynchronize(h->stream);
    cudaFree(t5); cudaFree(td); cudaFree(tc);
    if (e != cudaSuccess) { h->err = cudaGetErrorString(e); return MTD_E_CUDA; }
    return MTD_OK;
}

extern "C" int mtd_get_exposed(mtd_handle h, int32_t *agent_ids, int64_t cap, int64_t *n, int mem)
{
    if (!h || !n) return MTD_E_ARG;
    if (mem != MTD_MEM_HOST) { h->err = "mtd_get_exposed: host buffers only"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    StateCounters d;
    CK(cudaMemcpyAsync(&d, h->c.sctr, sizeof(d), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    int64_t cnt = std::min<int64_t>(d.n_applied, h->c.cap_expo);
    *n = cnt;
    if (cnt > cap) { h->err = "mtd_get_exposed: cap too small"; return MTD_E_CAPACITY; }
    std::vector<unsigned long long> tmp((size_t)cnt * 2);
    if (cnt) CK(cudaMemcpy(tmp.data(), h->expo_list, sizeof(unsigned long long) * 2 * (size_t)cnt, cudaMemcpyDeviceToHost));
    std::vector<std::pair<unsigned long long, int32_t>> v((size_t)cnt);
    for (int64_t i = 0; i < cnt; i++) v[(size_t)i] = { tmp[2 * i], (int32_t)tmp[2 * i + 1] };
    std::sort(v.begin(), v.end());          // order key = (cell order, row order): the reference's order of exposure
    for (int64_t i = 0; i < cnt; i++) agent_ids[i] = v[(size_t)i].second;
    return MTD_OK;
}

extern "C" int mtd_get_infection_prob(mtd_handle h, double *p, int64_t cap, int mem)
{
    // enables per-row probability capture on first use: call once with p == NULL before the day step to switch it on
    if (!h) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    if (!h->c.row_p) {
        double *rp; DA(rp, h->c.cap_rows); h->c.row_p = rp;
        if (!p) return MTD_OK;
        h->err = "probability capture was off during the last step"; return MTD_E_STATE;
    }
    if (!p) return MTD_OK;
    DevCounters d;
    CK(cudaMemcpyAsync(&d, h->c.ctr, sizeof(d), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (d.n_rows > cap) { h->err = "mtd_get_infection_prob: cap too small"; return MTD_E_CAPACITY; }
    CK(cudaMemcpyAsync(p, h->c.row_p, sizeof(double) * (size_t)d.n_rows, kind_out(mem), h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return MTD_OK;
}

extern "C" int mtd_occupancy(mtd_handle h, int64_t out[6])
{
    if (!h || !out) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    launch_occupancy(h->c, h->occ, h->stream, h->sm_count);
    unsigned long long t[6];
    CK(cudaMemcpyAsync(t, h->occ, sizeof(t), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    for (int k = 0; k < 6; k++) out[k] = (int64_t)t[k];
    return MTD_OK;
}

// ------------------------------------------------------------------------------------------------ K5
#include "k5_ctx.h"
int launch_k5(const TraceCtx &c, int32_t *flag_scan, long long *scan_tmp, cudaStream_t st, int sm_count);
int launch_k5_scan(const TraceCtx &c, cudaStream_t st, int sm_count);
int launch_k5_finish(const TraceCtx &c, int32_t *flag_scan, long long *scan_tmp, cudaStream_t st, int sm_count);

template <typename T>
static int regrow(mtd_handle h, T **p, int64_t *cap, int64_t need)
{
    if (need <= *cap) return 0;
    if (*p) cudaFree(*p);
    *p = nullptr; *cap = 0;
    void *q = nullptr;
    int64_t c = need + need / 2 + 1024;
    cudaError_t e = cudaMalloc(&q, (size_t)c * sizeof(T));
    if (e != cudaSuccess) { h->err = std::string("cudaMalloc: ") + cudaGetErrorString(e); return MTD_E_CUDA; }
    *p = (T *)q; *cap = c;
    return 0;
}

extern "C" int mtd_set_residences(mtd_handle h, int64_t n, const int64_t *res_off, const int32_t *res_nres,
                                  const int32_t *res_members, int mem)
{
    if (!h || n <= 0 || !res_off || !res_nres) { if (h) h->err = "mtd_set_residences: bad arguments"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    int64_t total = 0;
    if (mem == MTD_MEM_HOST) total = res_off[n];
    else CK(cudaMemcpy(&total, res_off + n, sizeof(int64_t), cudaMemcpyDeviceToHost));
    CK(cudaStreamSynchronize(h->stream));
    for (void *p : { (void *)h->res_off, (void *)h->res_nres, (void *)h->res_members, (void *)h->trace_sel,
                     (void *)h->res_first, (void *)h->res_k }) if (p) cudaFree(p);
    CK(cudaMalloc((void **)&h->res_off, sizeof(int64_t) * (size_t)(n + 1)));
    CK(cudaMalloc((void **)&h->res_nres, sizeof(int32_t) * (size_t)n));
    CK(cudaMalloc((void **)&h->res_members, sizeof(int32_t) * (size_t)std::max<int64_t>(total, 1)));
    CK(cudaMalloc((void **)&h->trace_sel, sizeof(int32_t) * (size_t)std::max<int64_t>(total, 1)));
    CK(cudaMalloc((void **)&h->res_first, sizeof(int32_t) * (size_t)n));
    CK(cudaMalloc((void **)&h->res_k, sizeof(int32_t) * (size_t)n));
    COPY_IN(h->res_off, res_off, sizeof(int64_t) * (size_t)(n + 1));
    COPY_IN(h->res_nres, res_nres, sizeof(int32_t) * (size_t)n);
    if (total) COPY_IN(h->res_members, res_members, sizeof(int32_t) * (size_t)total);
    CK(cudaStreamSynchronize(h->stream));
    h->n_res_cells = n; h->n_res_members = total;
    return MTD_OK;
}

static int trace_scratch(mtd_handle h)
{
    if (h->primary) return 0;
    const size_t N = (size_t)h->c.n_agents;
    CK(cudaMalloc((void **)&h->is_traced, N)); CK(cudaMalloc((void **)&h->primary, N)); CK(cudaMalloc((void **)&h->secondary, N));
    CK(cudaMalloc((void **)&h->d_epi_present, N));
    CK(cudaMalloc((void **)&h->trace_order, 4 * N)); CK(cudaMalloc((void **)&h->trace_cursor, 4 * N));
    CK(cudaMalloc((void **)&h->sec_owner, 4 * N)); CK(cudaMalloc((void **)&h->trace_scalars, 64));
    CK(cudaMalloc((void **)&h->trace_flagscan, 4 * (N + 2)));
    CK(cudaMalloc((void **)&h->trace_scan_tmp, sizeof(long long) * ((N + 2) / 2048 + 16)));
    return 0;
}

template <typename T>
static int trace_up(mtd_handle h, const T **dst, const T *src, size_t n)
{
    void *q = nullptr;
    cudaError_t e = cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T));
    if (e != cudaSuccess) { h->err = std::string("cudaMalloc(trace tape): ") + cudaGetErrorString(e); return MTD_E_CUDA; }
    h->trace_tape_allocs.push_back(q);
    if (n) cudaMemcpyAsync(q, src, n * sizeof(T), cudaMemcpyHostToDevice, h->stream);
    *dst = (const T *)q;
    return 0;
}

static int trace_prepare(mtd_handle h, const mtd_day *day, const mtd_trace_tape *tape, int64_t n_traced,
                         const int32_t *traced, int64_t n_slices, const int64_t *slice_off,
                         const int32_t *rows3, const uint8_t *epi_present, int mem, TraceCtx &t)
{
    if (!h || !day || n_traced < 0 || n_slices < 0 || n_slices > 8 || (n_slices && !slice_off)) {
        if (h) h->err = "mtd_contact_tracing: bad arguments"; return MTD_E_ARG; }
    if (!h->have_params || !h->res_off) { h->err = "mtd_contact_tracing: params / residences not set"; return MTD_E_STATE; }
    if (day->mode == MTD_MODE_REPLAY && !tape) { h->err = "replay mode without a trace tape"; return MTD_E_STATE; }
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    int r = trace_scratch(h);
    if (r) return r;
    CK(cudaStreamSynchronize(h->stream));
    for (void *p : h->trace_tape_allocs) cudaFree(p);
    h->trace_tape_allocs.clear();
    memset(&t, 0, sizeof(t));
    t.n_agents = h->c.n_agents; t.n_res_cells = h->n_res_cells; t.day = *day; t.P = h->d_params;
    if (tape && day->mode == MTD_MODE_REPLAY) {
        t.n_scan = tape->n_scan; t.n_prim = tape->n_prim;
        if ((r = trace_up(h, &t.scan_key, tape->scan_key, (size_t)tape->n_scan))) return r;
        if ((r = trace_up(h, &t.scan_u, tape->scan_u, (size_t)tape->n_scan))) return r;
        if ((r = trace_up(h, &t.prim_id, tape->prim_id, (size_t)tape->n_prim))) return r;
        if ((r = trace_up(h, &t.prim_off, tape->prim_off, (size_t)tape->n_prim + 1))) return r;
        size_t nd = tape->n_prim ? (size_t)tape->prim_off[tape->n_prim] : 0;
        if ((r = trace_up(h, &t.prim_draws, tape->prim_draws, nd))) return r;
    }
    const int64_t n_rows = n_slices ? slice_off[n_slices] : 0;
    t.n_traced = n_traced; t.n_slices = (int)n_slices;
    for (int64_t s = 0; s <= n_slices; s++) t.slice_off[s] = slice_off[s];
    if (mem == MTD_MEM_HOST) {
        if ((r = regrow(h, &h->d_traced, &h->cap_traced, n_traced))) return r;
        if ((r = regrow(h, &h->d_rows3, &h->cap_rows3, 3 * n_rows))) return r;
        if (n_traced) CK(cudaMemcpyAsync(h->d_traced, traced, 4 * (size_t)n_traced, cudaMemcpyHostToDevice, h->stream));
        if (n_rows) CK(cudaMemcpyAsync(h->d_rows3, rows3, 12 * (size_t)n_rows, cudaMemcpyHostToDevice, h->stream));
        t.traced = h->d_traced; t.rows3 = h->d_rows3;
        if (epi_present) { CK(cudaMemcpyAsync(h->d_epi_present, epi_present, (size_t)t.n_agents, cudaMemcpyHostToDevice, h->stream));
                           t.epi_present = h->d_epi_present; }
    } else { t.traced = traced; t.rows3 = rows3; t.epi_present = epi_present; }
    t.res_off = h->res_off; t.res_nres = h->res_nres; t.res_members = h->res_members; t.res_cellid = h->c.res_cellid;
    t.quarantine = h->c.quarantine; t.test_day = h->c.test_day; t.test_result_day = h->c.test_result_day;
    t.is_traced = h->is_traced; t.primary = h->primary; t.secondary = h->secondary;
    t.order = h->trace_order; t.n_order = h->trace_scalars + 8; t.cursor = h->trace_cursor;
    t.res_first = h->res_first; t.res_k = h->res_k; t.sel = h->trace_sel; t.sec_owner = h->sec_owner;
    t.stats = h->trace_scalars;
    return MTD_OK;
}

extern "C" int mtd_contact_tracing(mtd_handle h, const mtd_day *day, const mtd_trace_tape *tape, int64_t n_traced,
                                   const int32_t *traced, int64_t n_slices, const int64_t *slice_off,
                                   const int32_t *rows3, const uint8_t *epi_present, int mem)
{
    TraceCtx t;
    { int r = trace_prepare(h, day, tape, n_traced, traced, n_slices, slice_off, rows3, epi_present, mem, t); if (r) return r; }
    launch_k5(t, h->trace_flagscan, h->trace_scan_tmp, h->stream, h->sm_count);
    CK(cudaGetLastError());
    return MTD_OK;
}

// Cell-sharded contact tracing ("after K5: allgather of primary-contact ids; contact rows stay
// sharded"): every rank scans the contact rows of its own cells, the primary-contact flags are OR-ed over the ranks by
// the caller (one all-reduce of N bytes), and every rank then runs the scheduling phases on its replica of the records.
extern "C" int mtd_contact_tracing_scan(mtd_handle h, const mtd_day *day, int64_t n_traced, const int32_t *traced,
                                        int64_t n_slices, const int64_t *slice_off, const int32_t *rows3,
                                        const uint8_t *epi_present, uint8_t *primary_out, int mem)
{
    if (!h || !primary_out) { if (h) h->err = "mtd_contact_tracing_scan: bad arguments"; return MTD_E_ARG; }
    if (day && day->mode == MTD_MODE_REPLAY) { h->err = "mtd_contact_tracing_scan: Philox mode only (the replay tape is keyed by global row index)"; return MTD_E_ARG; }
    if (!h->trace_ctx) h->trace_ctx = new (std::nothrow) TraceCtx();
    if (!h->trace_ctx) return MTD_E_ARG;
    TraceCtx &t = *h->trace_ctx;
    { int r = trace_prepare(h, day, nullptr, n_traced, traced, n_slices, slice_off, rows3, epi_present, mem, t); if (r) return r; }
    launch_k5_scan(t, h->stream, h->sm_count);
    CK(cudaMemcpyAsync(primary_out, h->primary, (size_t)h->c.n_agents, cudaMemcpyDeviceToDevice, h->stream));
    CK(cudaGetLastError());
    h->trace_scan_pending = true;
    return MTD_OK;
}

extern "C" int mtd_contact_tracing_finish(mtd_handle h, const uint8_t *primary_in)
{
    if (!h || !primary_in) { if (h) h->err = "mtd_contact_tracing_finish: bad arguments"; return MTD_E_ARG; }
    if (!h->trace_scan_pending) { h->err = "mtd_contact_tracing_finish before mtd_contact_tracing_scan"; return MTD_E_STATE; }
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    CK(cudaMemcpyAsync(h->primary, primary_in, (size_t)h->c.n_agents, cudaMemcpyDeviceToDevice, h->stream));
    launch_k5_finish(*h->trace_ctx, h->trace_flagscan, h->trace_scan_tmp, h->stream, h->sm_count);
    CK(cudaGetLastError());
    h->trace_scan_pending = false;
    return MTD_OK;
}

int launch_k6(const DevCtx &c, const uint8_t *epi_present, uint8_t *positive, int32_t *flag_scan, long long *scan_tmp,
              int32_t *ids, int64_t cap, int32_t *n_out, int64_t rp_n, const int32_t *rp_agent, const double *rp_u,
              cudaStream_t st, int sm_count);

extern "C" int mtd_set_k6_tape(mtd_handle h, int64_t n, const int32_t *agent_ids, const double *draws2)
{
    if (!h || n < 0 || (n && (!agent_ids || !draws2))) { if (h) h->err = "mtd_set_k6_tape: bad arguments"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    if (h->k6_agent) { cudaFree(h->k6_agent); h->k6_agent = nullptr; }
    if (h->k6_u) { cudaFree(h->k6_u); h->k6_u = nullptr; }
    h->k6_n = 0;
    if (n == 0) return MTD_OK;
    for (int64_t i = 1; i < n; i++) if (agent_ids[i] <= agent_ids[i - 1]) { h->err = "mtd_set_k6_tape: agent ids must be strictly ascending"; return MTD_E_ARG; }
    CK(cudaMalloc((void **)&h->k6_agent, 4 * (size_t)n)); CK(cudaMalloc((void **)&h->k6_u, 16 * (size_t)n));
    CK(cudaMemcpyAsync(h->k6_agent, agent_ids, 4 * (size_t)n, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->k6_u, draws2, 16 * (size_t)n, cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    h->k6_n = n;
    return MTD_OK;
}

extern "C" int mtd_process_test_results(mtd_handle h, const mtd_day *day, const uint8_t *epi_present, int32_t *traced_ids,
                                        int64_t cap, int64_t *n, int mem)
{
    if (!h || !day || !n || cap < 0 || (cap && !traced_ids)) { if (h) h->err = "mtd_process_test_results: bad arguments"; return MTD_E_ARG; }
    if (!h->have_params) { h->err = "mtd_process_test_results: params not set"; return MTD_E_STATE; }
    CK(cudaSetDevice(h->cfg.device));
    TOUCH();
    { int r = trace_scratch(h); if (r) return r; }
    h->c.day = *day;
    const size_t N = (size_t)h->c.n_agents;
    const uint8_t *present = epi_present;
    if (epi_present && mem == MTD_MEM_HOST) {
        CK(cudaMemcpyAsync(h->d_epi_present, epi_present, N, cudaMemcpyHostToDevice, h->stream));
        present = h->d_epi_present;
    }
    { int r = regrow(h, &h->d_traced, &h->cap_traced, (int64_t)N); if (r) return r; }
    if (day->mode == MTD_MODE_REPLAY && h->k6_n == 0) { h->err = "mtd_process_test_results: replay mode without mtd_set_k6_tape"; return MTD_E_STATE; }
    launch_k6(h->c, present, h->is_traced, h->trace_flagscan, h->trace_scan_tmp, h->d_traced, (int64_t)N, h->trace_scalars + 12,
              h->k6_n, h->k6_agent, h->k6_u, h->stream, h->sm_count);
    int32_t cnt = 0;
    CK(cudaMemcpyAsync(&cnt, h->trace_scalars + 12, sizeof(cnt), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    *n = cnt;
    if (cnt > cap) { h->err = "mtd_process_test_results: cap too small"; return MTD_E_CAPACITY; }
    if (cnt) CK(cudaMemcpyAsync(traced_ids, h->d_traced, 4 * (size_t)cnt, kind_out(mem), h->stream));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    return MTD_OK;
}

extern "C" int mtd_get_traced(mtd_handle h, uint8_t *primary, uint8_t *secondary, mtd_trace_stats *stats, int mem)
{
    if (!h || !h->primary) { if (h) h->err = "mtd_get_traced before mtd_contact_tracing"; return MTD_E_STATE; }
    CK(cudaSetDevice(h->cfg.device));
    const size_t N = (size_t)h->c.n_agents;
    COPY_OUT(primary, h->primary, N); COPY_OUT(secondary, h->secondary, N);
    int32_t s[4];
    CK(cudaMemcpyAsync(s, h->trace_scalars, sizeof(s), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (stats) { stats->n_scan_draws = s[0]; stats->tape_miss = s[1]; stats->n_primary = s[2]; stats->n_secondary = s[3]; }
    return MTD_OK;
}


// ------------------------------------------------------------------------------------------------ pipelined boundary
static int pipeline_init(mtd_handle h)
{
    if (h->copy_stream) return MTD_OK;
    int lo = 0, hi = 0;
    CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CK(cudaStreamCreateWithPriority(&h->copy_stream, cudaStreamNonBlocking, hi));
    CK(cudaStreamCreateWithPriority(&h->export_stream, cudaStreamNonBlocking, hi));
    CK(cudaStreamCreateWithPriority(&h->bulk_stream, cudaStreamNonBlocking, hi));
    for (cudaEvent_t *e : { &h->ev_staged[0], &h->ev_staged[1], &h->ev_k0_done, &h->ev_snap, &h->ev_exported[0], &h->ev_exported[1],
                            &h->ev_res[0], &h->ev_res[1] })
        CK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    return MTD_OK;
}

extern "C" int mtd_set_export(mtd_handle h, int on)
{
    if (!h) return MTD_E_ARG;
    CK(cudaSetDevice(h->cfg.device));
    if (on && !h->rows12_buf[0]) {
        int r = pipeline_init(h); if (r) return r;
        for (int k = 0; k < 2; k++) {
            DA(h->rows12_buf[k], 3 * h->c.cap_rows + 4);
            DA(h->snap[k], sizeof(mtd_day_result) / sizeof(int64_t));
            DA(h->expo_snap[k], 2 * h->c.cap_expo + 2);
        }
    }
    h->export_on = on != 0;
    return MTD_OK;
}

extern "C" int mtd_prefetch_intervals(mtd_handle h, int64_t M, const int32_t *iv_agent, const int32_t *iv_cell,
                                      const uint8_t *iv_start, const uint8_t *iv_end, const uint32_t *iv_order)
{
    if (!h || M < 0 || !iv_agent || !iv_cell || !iv_start || !iv_end) return MTD_E_ARG;
    if (M > h->c.cap_intervals) { h->err = "interval stream exceeds max_intervals"; return MTD_E_CAPACITY; }
    CK(cudaSetDevice(h->cfg.device));
    { int r = pipeline_init(h); if (r) return r; }
    const int sl = h->st_next;
    if (!h->st_agent[sl]) {
        const int64_t cap = h->c.cap_intervals;
        DA(h->st_agent[sl], cap); DA(h->st_cell[sl], cap); DA(h->st_start[sl], cap); DA(h->st_end[sl], cap); DA(h->st_order[sl], cap);
    }
    // this slot was last read by the K0 of two days ago; one event after the most recent K0 covers it
    CK(cudaStreamWaitEvent(h->copy_stream, h->ev_k0_done, 0));
    const size_t m = (size_t)M;
    CK(cudaMemcpyAsync(h->st_agent[sl], iv_agent, 4 * m, cudaMemcpyHostToDevice, h->copy_stream));
    CK(cudaMemcpyAsync(h->st_cell[sl], iv_cell, 4 * m, cudaMemcpyHostToDevice, h->copy_stream));
    CK(cudaMemcpyAsync(h->st_start[sl], iv_start, m, cudaMemcpyHostToDevice, h->copy_stream));
    CK(cudaMemcpyAsync(h->st_end[sl], iv_end, m, cudaMemcpyHostToDevice, h->copy_stream));
    if (iv_order) CK(cudaMemcpyAsync(h->st_order[sl], iv_order, 4 * m, cudaMemcpyHostToDevice, h->copy_stream));
    CK(cudaEventRecord(h->ev_staged[sl], h->copy_stream));
    h->st_n[sl] = M; h->st_has_order[sl] = iv_order != nullptr;
    h->st_ready = sl; h->st_next = sl ^ 1;
    return MTD_OK;
}

// snapshot of what the next day step overwrites, in two halves: the contact counters (cleared by the next K0 on the main
// stream) and the state chain's products (exposure list, applied count, occupancy; state stream, after the apply)
__global__ void export_snap_contact_kernel(DevCtx c, int64_t *snap)
{
    snap[0] = c.day.day; snap[1] = c.ctr->n_rows; snap[9] = (int64_t)c.ctr->pairs_evaluated; snap[10] = c.ctr->error;
}

__global__ void export_snap_state_kernel(DevCtx c, const unsigned long long *occ, const unsigned long long *expo_list,
                                         int64_t *snap, unsigned long long *expo_snap, int64_t ticket)
{
    const int n = c.sctr->n_applied < (int)c.cap_expo ? c.sctr->n_applied : (int)c.cap_expo;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * n; i += gridDim.x * blockDim.x) expo_snap[i] = expo_list[i];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        snap[2] = n;
        for (int k = 0; k < 6; k++) snap[3 + k] = (int64_t)occ[k];
        if (snap[10] == 0) snap[10] = c.sctr->error;
        snap[11] = ticket;
    }
}

// Host exports, second half: once the day's result record has landed (ev_res) the row and exposure counts are known on
// the host, and the two bulk copies are issued with their exact sizes on the bulk stream -- copy-engine transfers
// that leave every SM to the next day's kernels.
static int issue_host_export(mtd_handle h, int sl)
{
    mtd_handle_s::HostExport &x = h->hx[sl];
    if (!x.unissued) return MTD_OK;
    x.unissued = false;
    CK(cudaEventSynchronize(h->ev_res[sl]));
    const int64_t n_rows = std::min<int64_t>(x.result->n_rows, x.cap_rows);
    const int64_t n_exp = std::min<int64_t>(x.result->n_exposed, x.cap_exposed);
    // the bulk copies have a stream of their own: the export stream already holds the NEXT day's result record behind a
    // wait for that day's snapshot, and a copy queued there would start a whole day late
    if (x.rows && n_rows > 0)
        CK(cudaMemcpyAsync(x.rows, h->rows12_buf[sl], 12 * (size_t)n_rows, cudaMemcpyDeviceToHost, h->bulk_stream));
    if (x.kv && n_exp > 0)
        CK(cudaMemcpyAsync(x.kv, h->expo_snap[sl], 16 * (size_t)n_exp, cudaMemcpyDeviceToHost, h->bulk_stream));
    CK(cudaEventRecord(h->ev_exported[sl], h->bulk_stream));
    return MTD_OK;
}

extern "C" int mtd_export_day_async(mtd_handle h, void *rows12, int64_t cap_rows, uint64_t *exposed_kv, int64_t cap_exposed,
                                    mtd_day_result *result, int mem, int64_t *ticket)
{
    if (!h || !result) return MTD_E_ARG;
    if (!h->export_on || !h->day_ready || !h->c.rows12) { h->err = "mtd_export_day_async: mtd_set_export(h, 1) before the day step"; return MTD_E_STATE; }
    CK(cudaSetDevice(h->cfg.device));
    const int sl = (int)((h->day_seq - 1) & 1);          // slot of the day just computed
    { int r = issue_host_export(h, sl); if (r) return r; }
    const int64_t tk = ++h->export_ticket;
    // snapshot of everything the next day step overwrites: contact counters on the main stream (before the next K0),
    // exposure list + occupancy on the state stream (after the apply)
    export_snap_contact_kernel<<<1, 1, 0, h->stream>>>(h->c, h->snap[sl]);
    { int r = state_after_main(h); if (r) return r; }
    launch_occupancy(h->c, h->occ, h->state_stream, h->sm_count);
    export_snap_state_kernel<<<32, 256, 0, h->state_stream>>>(h->c, h->occ, h->expo_list, h->snap[sl], h->expo_snap[sl], tk);
    h->state_dirty = true;
    CK(cudaEventRecord(h->ev_snap, h->state_stream));
    CK(cudaStreamWaitEvent(h->export_stream, h->ev_snap, 0));
    if (mem == MTD_MEM_HOST) {
        // the result record first; mtd_export_wait (or the next reuse of this slot) issues the bulk copies
        CK(cudaMemcpyAsync(result, h->snap[sl], sizeof(mtd_day_result), cudaMemcpyDeviceToHost, h->export_stream));
        CK(cudaEventRecord(h->ev_res[sl], h->export_stream));
        mtd_handle_s::HostExport &x = h->hx[sl];
        x.unissued = true; x.rows = rows12; x.cap_rows = cap_rows; x.kv = exposed_kv; x.cap_exposed = cap_exposed;
        x.result = result;
        h->launches[MTD_STAGE_K2] += 2;
    } else {
        // device destinations: counts are read on the device by a streaming kernel on the side stream
        if (rows12) launch_stream_out(rows12, h->rows12_buf[sl], h->snap[sl] + 1, 12, cap_rows, 2 * h->sm_count, h->export_stream);
        if (exposed_kv) launch_stream_out(exposed_kv, h->expo_snap[sl], h->snap[sl] + 2, 16, cap_exposed, 8, h->export_stream);
        CK(cudaMemcpyAsync(result, h->snap[sl], sizeof(mtd_day_result), cudaMemcpyDefault, h->export_stream));
        CK(cudaEventRecord(h->ev_exported[sl], h->export_stream));
        h->launches[MTD_STAGE_K2] += 2 + (rows12 ? 1 : 0) + (exposed_kv ? 1 : 0);
    }
    h->exported_pending[sl] = true;
    h->ticket_slot[tk & 1] = sl;
    CK(cudaGetLastError());
    if (ticket) *ticket = tk;
    return MTD_OK;
}

extern "C" int mtd_export_wait(mtd_handle h, int64_t ticket)
{
    if (!h) return MTD_E_ARG;
    if (ticket <= 0 || ticket > h->export_ticket || ticket + 2 <= h->export_ticket) { h->err = "mtd_export_wait: unknown ticket"; return MTD_E_ARG; }
    CK(cudaSetDevice(h->cfg.device));
    const int sl = h->ticket_slot[ticket & 1];
    { int r = issue_host_export(h, sl); if (r) return r; }
    CK(cudaEventSynchronize(h->ev_exported[sl]));
    return MTD_OK;
}
