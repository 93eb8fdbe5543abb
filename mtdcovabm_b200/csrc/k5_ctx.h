// kernel-argument block of K5 (contact tracing), shared by k5_tracing.cu and capi.cu
#pragma once
#include "common.cuh"

struct TraceCtx {
    int64_t n_agents, n_res_cells;
    mtd_day day;
    const mtd_params *P;
    // replay tape (device copies)
    int64_t n_scan; const uint64_t *scan_key; const double *scan_u;
    int64_t n_prim; const int32_t *prim_id; const int32_t *prim_off; const double *prim_draws;
    // inputs
    int64_t n_traced; const int32_t *traced;
    int n_slices; int64_t slice_off[9];
    const int32_t *rows3;
    const uint8_t *epi_present;
    const int64_t *res_off; const int32_t *res_nres; const int32_t *res_members;
    const int32_t *res_cellid;
    // state touched
    int32_t *quarantine, *test_day, *test_result_day;
    // scratch / outputs
    uint8_t *is_traced, *primary, *secondary;
    int32_t *order;          // [N] processing order of the primary contacts
    int32_t *n_order;        // device scalar
    int32_t *cursor;         // [N] per-order-index tape cursor
    int32_t *res_first;      // [n_res_cells] smallest processing index among the residence's primaries
    int32_t *res_k;          // [n_res_cells] number of selected secondary contacts
    int32_t *sel;            // [len(res_members)] selection scratch, segment of a residence at res_off
    int32_t *sec_owner;      // [N] smallest processing index among the residences that selected the agent
    int32_t *stats;          // [4] n_scan_draws, tape_miss, n_primary, n_secondary
};

