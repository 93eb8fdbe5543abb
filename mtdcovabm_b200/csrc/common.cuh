// mtdcovabm_b200/csrc/common.cuh -- shared device structures and helpers (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/mtd_b200.h"

#define MTD_NUM_CLASSES 9      // cells by interval count: <=8, <=16, <=32 (sub-warp groups); <=64, <=128, <=192, <=320, <=1024
                               //   (block per cell, shared memory), above (block per cell, global workspace)
#define MTD_LOGTAB_N 4096      // host-computed math.log(i) for small integers
#define FULLMASK 0xffffffffu

// enums of the reference (epidemiologyclasses.py, cellsclasses.py)
enum { S_UNDEF = 0, S_SUS = 1, S_EXP = 2, S_INF = 3, S_REC = 4, S_DEC = 5 };
enum { IT_UNDEF = 0, IT_PREASYMP = 1, IT_PRESYMP = 2, IT_ASYMP = 3, IT_SYMP = 4 };
enum { SV_UNDEF = 0, SV_MILD = 1, SV_SEVERE = 2, SV_CRITICAL = 3 };
enum { T_EXP_INF = 0, T_INF_SYMP = 1, T_SYMP_SEV = 2, T_SEV_CRI = 3, T_CRI_DEATH = 4, T_ASYMP_REC = 5,
       T_MILD_REC = 6, T_SEV_REC = 7, T_CRI_REC = 8, T_REC_SUS = 9 };
enum { CT_HOUSEHOLD = 0, CT_WORKPLACE = 1, CT_ACCOM = 2, CT_HOSPITAL = 3, CT_ENTERTAINMENT = 4, CT_SCHOOL = 5,
       CT_CLASSROOM = 6, CT_INSTITUTION = 7, CT_TRANSPORT = 8, CT_RELIGION = 9, CT_AIRPORT = 10 };
enum { SCT_RESIDENCE = 0, SCT_WORKPLACE = 1, SCT_SCHOOL = 2, SCT_COMMUNITY = 3 };
enum { SITE_SELECT = 1, SITE_PAIR = 2, SITE_EXPO = 3, SITE_TRACE = 4 };

// device-side counters of one day step (copied to mtd_counters by mtd_get_counters)
struct DevCounters {
    unsigned long long pairs_evaluated;
    unsigned long long n_potential;
    unsigned long long n_incidences;    // sum over cells of distinct agents (the U of the algorithmic-bytes formula)
    unsigned long long row_bound_total; // total scratch rows requested
    int n_slots;
    int n_rows;
    int n_succ;                          // successful S-I draws (before winner resolution)
    int n_delta;                         // delta records produced locally
    int n_applied_unused;
    int n_si_rows;
    int sched_overflow;
    int tape_miss;
    int class_cells[MTD_NUM_CLASSES];
    int class_cursor[MTD_NUM_CLASSES];   // work-fetch cursors of the K1 blocks
    int class_started[MTD_NUM_CLASSES];  // K1 blocks of the class that have started (slot turnover, k1_block.cu)
    int class_cursor2[MTD_NUM_CLASSES];  // work-fetch cursors of the thread-per-cell chain kernels (k1_chain.cu)
    int error;
    int pad;
};

// counters of the state chain (apply of the day's exposure records), which runs on the handle's state stream beside the
// next day's K0/K1/K2: kept apart from DevCounters, which the next day's K0 clears
struct StateCounters {
    int n_applied;                       // exposures applied
    int sched_overflow;
    int error;                           // MTD_E_TIMEOUT of the peer wait
    int pad;
};

// multi-GPU exchange of the exposure records over NVLink peer memory (one slice per source rank and day parity)
#define MTD_MAX_WORLD 16
struct DevExchange {
    int rank, world;
    int64_t cap;                                   // records per slice (+1 header record)
    int32_t *buf[MTD_MAX_WORLD];                   // base of rank r's exchange buffer [2][world][cap+1][MTD_DELTA_WORDS] (peer mapped)
    unsigned long long *flags[MTD_MAX_WORLD];      // rank r's arrival flags [2][world]: sequence number of the last slice landed
};

// one CSR record of the cell -> interval layout (16 B, one vector load)
struct __align__(16) IvRec {
    int32_t agent;
    uint32_t order;      // emission index (position in the day's stream)
    uint32_t se;         // start | end << 8
    uint32_t pad;
};

// one dense contact row (16 B)
struct __align__(16) RowRec {
    int32_t a1, a2;
    uint32_t sed;        // start | end << 8 | duration << 16
    int32_t slot;        // active-cell slot (cell id = slot_cell[slot])
};

struct DevTape {
    int64_t n_sel;  const uint64_t *sel_key;  const int32_t *sel_off;  const int32_t *sel_picks;
    int64_t n_pair; const uint64_t *pair_k1;  const uint32_t *pair_k2; const double *pair_u;
    int64_t n_expo; const int32_t *expo_agent; const double *expo_row;
};

// everything a kernel needs, passed by value
// the two static attributes K1 reads per distinct agent of a cell (contactnetwork.py:237-252), packed by mtd_set_static
struct __align__(16) AgentAttr { double soc_rate; int32_t age_bracket; int32_t pad; };

struct DevCtx {
    // capacities
    int64_t n_agents, n_cells, cap_intervals, cap_rows, cap_scratch_rows, cap_expo;
    // day
    mtd_day day;
    DevTape tape;
    // static
    const int8_t *cell_type; const int16_t *cell_indid;
    const int8_t *age_bracket; const int8_t *epi_age_bracket; const double *soc_rate; const int32_t *res_cellid;
    const AgentAttr *k1_attr;  // {soc_rate, age_bracket} of every agent in ONE 16-byte record: K1's gather touches one sector
    // state
    int8_t *seir; int8_t *inf_type; int8_t *inf_sev; int32_t *last_dose_day;
    int16_t *sched_next;       // [n_agents] earliest day among the agent's schedule entries (MTD_SCHED_NEXT_NONE = none): K4's filter
    int32_t *sched; uint8_t *sched_cnt; int32_t *test_day; int32_t *test_result_day; int32_t *quarantine; int32_t *hosp;
    // day inputs
    int64_t M;
    const int32_t *iv_agent; const int32_t *iv_cell; const uint8_t *iv_start; const uint8_t *iv_end;
    const uint32_t *iv_order;
    // K0 products
    int32_t *cell_count;       // [n_cells]
    uint32_t *cell_first;      // [n_cells] emission index of first record
    int64_t *cell_base;        // [n_cells] CSR offset of the cell's segment
    uint16_t *iv_pos;          // [M] position of the record inside its cell segment (arrival order)
    int32_t *flag_scan;        // [M+64] scanned per-word counts of the "first record of its cell" flags + the flag words
    int32_t *slot_cell;        // [slots]
    int32_t *slot_m;           // [slots]
    uint32_t *slot_first;      // [slots] order key of the cell
    uint8_t *slot_act;         // [slots] activity column of the cell (contactnetwork.py:317-347), resolved once by K0
    int64_t *slot_off;         // [slots+1] CSR offsets
    int64_t *slot_rowoff;      // [slots+1] scratch-row offsets
    int32_t *slot_rowbound;    // [slots]
    int32_t *slot_nrows;       // [slots] rows produced
    int64_t *slot_denseoff;    // [slots+1]
    int32_t *class_list;       // [MTD_NUM_CLASSES + 7][slots cap]; the extra regions hold classes 6, 5, 0, 1, 2, 4, 3 sorted by cell size (K1)
    int64_t class_stride;
    IvRec *csr;                // [M]
    // K1 products
    RowRec *scratch_rows;      // [cap_scratch_rows]
    uint32_t *rows12;          // [3 * cap_rows] packed 12-byte rows of the day for the async export (NULL = off)
    int32_t *k1_bins;          // [2][3][32] histogram and cursors of the small-class size sort (K1)
    uint32_t *k1_bits;         // [k1_bits_words] per-warp adjacency bit matrices of the legacy warp-per-cell K1 (k1_legacy.cu)
    int64_t k1_bits_words;
    unsigned char *k1_rec;     // [80 * cap_intervals] per-cell chain records of the split K1 (K1Rec below), at 80 * slot_off[slot]
    int32_t k1_split;          // the day's cells of 33..320 intervals go through prepare / chain / row-build kernels
    int32_t max_cell_m;        // largest cell (intervals) this handle takes: min(MTD_MAX_CELL_INTERVALS, max_intervals)
    RowRec *rows;              // [cap_rows] dense
    // K3
    unsigned long long *win;   // [n_agents] winner key per agent
    int32_t *succ;             // [cap_rows] dense row indices of successful draws
    double *row_p;             // [cap_rows] optional infection probability per row (NULL = off)
    int32_t *delta;            // [cap_expo][MTD_DELTA_WORDS]
    DevCounters *ctr;
    int32_t *n_succ_pub;       // the day's n_succ, published by K3 for the apply chain (state stream), which must not
                               //   read the day's counter block: the K0 of two days later clears it on another stream
    StateCounters *sctr;
    const mtd_params *P;       // device copy of the parameters
    // intervention mask (mtd_set_intervention_mask): per-agent "stays in the residence cell today" flag, or NULL = off
    uint8_t *agent_masked;
    // caller-supplied contacts (mtd_set_contacts): overlap lists per dense row instead of the cell's interval records
    const int64_t *ov_off;     // [n_rows + 1] or NULL
    const uint8_t *ov_se;      // [2 * total] start, end per overlap
    const double *logtab;      // [MTD_LOGTAB_N] glibc log(i)
};


// ---------------------------------------------------------------- split K1: the per-cell record between its three kernels
// Cells of 33..320 intervals: k1_block.cu (a block per cell) prepares the cell -- agents, adjacency rows, degrees, greedy
// order -- and leaves this record in HBM at k1_rec + 80 * slot_off[slot] (80 bytes per interval of the cell: enough for
// every cell of up to 336 intervals); k1_chain.cu runs the sequential greedy sampling with a GROUP OF 8 LANES per cell on it (4
// cells per warp); K2 builds the contact rows of the selected pairs from it.
//   header  int4 {n_act, n, m, bwm}; alive0 u32[bw2] (first positions of the agents with degree > 0)
//   mains   n_act records in greedy order, (6 + bw2) words each: {x | first position << 16, agent id, the agent's first
//           Philox block (4 draws), adjacency row[bw2]}
//   agent   i32[n] agent id per local agent;  deg u16[m] initial degree at an agent's first position (0 elsewhere)
//   se u16[m] start | end << 8;  la u16[m] local agent of a position;  apos u16[m], aoff u16[n + 1] interval lists
struct K1Rec {
    uint32_t bw2, main_words, o_alive, o_main, o_agent, o_deg, o_se, o_la, o_apos, o_aoff;
    __host__ __device__ K1Rec(int m, int n, int bwm)
    {
        bw2 = (uint32_t)((bwm + 1) & ~1); main_words = 6 + bw2;
        uint32_t o = 16;
        o_alive = o; o += 48;
        o_main = o;  o += (uint32_t)n * main_words * 4; o = (o + 15u) & ~15u;
        o_agent = o; o += 4u * (uint32_t)n; o = (o + 15u) & ~15u;
        o_deg = o;   o += 2u * (uint32_t)m; o = (o + 15u) & ~15u;
        o_se = o;    o += 2u * (uint32_t)m; o = (o + 15u) & ~15u;
        o_la = o;    o += 2u * (uint32_t)m; o = (o + 15u) & ~15u;
        o_apos = o;  o += 2u * (uint32_t)m; o = (o + 15u) & ~15u;
        o_aoff = o;
    }
};
#define MTD_SCHED_NEXT_NONE 0x7f7f   // "no entry" of DevCtx::sched_next (what a byte-wise memset of 0x7f leaves; above every day)
#define MTD_K1_REC_STRIDE 80       // bytes of record space per interval
#define MTD_K1_SPLIT_MAX_M 320     // largest cell the split path takes

// streams / events of the split K1's chain kernels (owned by the handle): the chain kernel of a class waits for the class's
// prepare kernel and runs beside the prepare kernels of the next classes; the stage joins on `done`
// Development aid (MTD_K1_TIMELINE=<n>): the n-th K1 stage of the process records an event behind every launch on the
// launch's stream; the next K1 stage prints, per mark, when the work ahead of it on its stream had ended (k1_contacts.cu).
void k1_tl_mark(const char *name, cudaStream_t st);
// st[k] / st2[k]: chain stream of class 3 + k for days of even / odd parity; pst[k]: the class's prepare stream; par: parity
// of the day being launched
struct K1ChainStreams { cudaStream_t st[4]; cudaEvent_t prep[4], done[4]; cudaEvent_t fork; cudaStream_t st2[4], pst[4]; int par; };

// ---------------------------------------------------------------- Philox4x32-10
__host__ __device__ __forceinline__ void philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                    uint32_t k1, uint32_t out[4])
{
#pragma unroll
    for (int r = 0; r < 10; r++) {
#ifdef __CUDA_ARCH__
        uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
#else
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t h0 = (uint32_t)(p0 >> 32), l0 = (uint32_t)p0, h1 = (uint32_t)(p1 >> 32), l1 = (uint32_t)p1;
#endif
        uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// w[i] without a dynamically indexed local array
__device__ __forceinline__ uint32_t pick4(const uint32_t w[4], int i)
{
    return i == 0 ? w[0] : (i == 1 ? w[1] : (i == 2 ? w[2] : w[3]));
}

__device__ __forceinline__ double u53(uint32_t a, uint32_t b)
{
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) / 9007199254740992.0;
}

// contactnetwork.py:317-347
__device__ __forceinline__ int activity_of_celltype(int ct, int indid)
{
    if (indid == 9 && ct != CT_ACCOM) return 5;
    switch (ct) {
    case CT_HOUSEHOLD: case CT_ACCOM: case CT_INSTITUTION: return 1;
    case CT_WORKPLACE: return 3;
    case CT_HOSPITAL: case CT_ENTERTAINMENT: case CT_RELIGION: case CT_AIRPORT: return 5;
    case CT_SCHOOL: return 2;
    case CT_TRANSPORT: return 4;
    default: return 0;   // Classroom -> "total" column (reference quirk)
    }
}

// util.py:239-268
__host__ __device__ __forceinline__ int simcelltype_of(int ct)
{
    switch (ct) {
    case CT_HOUSEHOLD: case CT_ACCOM: case CT_INSTITUTION: return SCT_RESIDENCE;
    case CT_WORKPLACE: return SCT_WORKPLACE;
    case CT_SCHOOL: case CT_CLASSROOM: return SCT_SCHOOL;
    default: return SCT_COMMUNITY;
    }
}

__device__ __forceinline__ uint32_t lanemask_lt()
{
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

