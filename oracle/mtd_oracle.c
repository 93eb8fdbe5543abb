/*
 * oracle/mtd_oracle.c -- CPU restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY (see mtd_oracle.h).
 * Sequential, single-threaded, written to mirror the reference's control flow one-to-one so that a reader can
 * check each block against the cited reference lines.  Build: see oracle/Makefile (gcc -O2 -shared -fPIC).
 */
#include "mtd_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ enums (epidemiologyclasses.py, cellsclasses.py) */
enum { S_UNDEF = 0, S_SUS = 1, S_EXP = 2, S_INF = 3, S_REC = 4, S_DEC = 5 };
enum { IT_UNDEF = 0, IT_PREASYMP = 1, IT_PRESYMP = 2, IT_ASYMP = 3, IT_SYMP = 4 };
enum { SV_UNDEF = 0, SV_MILD = 1, SV_SEVERE = 2, SV_CRITICAL = 3 };
enum { T_EXP_INF = 0, T_INF_SYMP = 1, T_SYMP_SEV = 2, T_SEV_CRI = 3, T_CRI_DEATH = 4, T_ASYMP_REC = 5,
       T_MILD_REC = 6, T_SEV_REC = 7, T_CRI_REC = 8, T_REC_SUS = 9 };
enum { CT_HOUSEHOLD = 0, CT_WORKPLACE = 1, CT_ACCOM = 2, CT_HOSPITAL = 3, CT_ENTERTAINMENT = 4, CT_SCHOOL = 5,
       CT_CLASSROOM = 6, CT_INSTITUTION = 7, CT_TRANSPORT = 8, CT_RELIGION = 9, CT_AIRPORT = 10 };
enum { SCT_RESIDENCE = 0, SCT_WORKPLACE = 1, SCT_SCHOOL = 2, SCT_COMMUNITY = 3 };
enum { Q_SYMPTOMATIC = 0, Q_POSITIVE = 1, Q_POSITIVE_CONTACT = 2, Q_SECONDARY_CONTACT = 3 };

/* ------------------------------------------------------------------ Philox4x32-10 (Salmon et al. 2011) */
void mtdo_philox4x32(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

double mtdo_u53(uint32_t a, uint32_t b)
{
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) / 9007199254740992.0;
}

static void philox_site(const mtdo_dayctx *D, uint32_t site, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t out[4])
{
    uint32_t ctr[4] = { (uint32_t)D->day, c1, c2, c3 }, key[2] = { D->seed, site };
    mtdo_philox4x32(ctr, key, out);
}

/* ------------------------------------------------------------------ numpy float64 add.reduce (pairwise, blocks of 128) */
double mtdo_np_pairwise_sum(const double *a, int64_t n)
{
    if (n < 8) {
        double r = -0.0;
        for (int64_t i = 0; i < n; i++) r += a[i];
        return r;
    } else if (n <= 128) {
        double r[8];
        int64_t i;
        for (int j = 0; j < 8; j++) r[j] = a[j];
        for (i = 8; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; j++) r[j] += a[i + j];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; i++) res += a[i];
        return res;
    } else {
        int64_t n2 = n / 2;
        n2 -= n2 % 8;
        return mtdo_np_pairwise_sum(a, n2) + mtdo_np_pairwise_sum(a + n2, n - n2);
    }
}

/* Python round() on a float: round-half-even, correctly rounded -> rint() in the default rounding mode */
static int64_t py_round(double x) { return (int64_t)rint(x); }

/* ------------------------------------------------------------------ small tables */
int mtdo_activity_of_celltype(int ct, int indid)   /* contactnetwork.py:317-347 */
{
    if (indid == 9 && ct != CT_ACCOM) return 5;
    switch (ct) {
    case CT_HOUSEHOLD: return 1;
    case CT_WORKPLACE: return 3;
    case CT_ACCOM: return 1;
    case CT_HOSPITAL: return 5;
    case CT_ENTERTAINMENT: return 5;
    case CT_SCHOOL: return 2;
    case CT_INSTITUTION: return 1;
    case CT_TRANSPORT: return 4;
    case CT_RELIGION: return 5;
    case CT_AIRPORT: return 5;
    default: return 0;   /* Classroom falls here (quirk B5) */
    }
}

int mtdo_simcelltype(int ct)   /* util.py:239-268 */
{
    switch (ct) {
    case CT_HOUSEHOLD: case CT_ACCOM: case CT_INSTITUTION: return SCT_RESIDENCE;
    case CT_WORKPLACE: return SCT_WORKPLACE;
    case CT_SCHOOL: case CT_CLASSROOM: return SCT_SCHOOL;
    default: return SCT_COMMUNITY;
    }
}

void mtdo_transition(int seir, int it, int sv, int trans, int *ns, int *nt, int *nv)   /* seirstateutil.py:105-136 */
{
    switch (trans) {
    case T_EXP_INF:  *ns = S_INF; *nt = (it == IT_PREASYMP) ? IT_ASYMP : it; *nv = SV_MILD; return;
    case T_INF_SYMP: *ns = IT_SYMP /* quirk B1: InfectionType.Symptomatic (4) lands in the SEIR slot */;
                     *nt = (it == IT_PRESYMP) ? IT_SYMP : it; *nv = sv; return;
    case T_SYMP_SEV: *ns = seir; *nt = it; *nv = SV_SEVERE; return;
    case T_SEV_CRI:  *ns = seir; *nt = it; *nv = SV_CRITICAL; return;
    case T_CRI_DEATH: *ns = S_DEC; *nt = it; *nv = sv; return;
    case T_ASYMP_REC: case T_MILD_REC: case T_SEV_REC: case T_CRI_REC:
                     *ns = S_REC; *nt = IT_UNDEF; *nv = SV_UNDEF; return;
    case T_REC_SUS:  *ns = S_SUS; *nt = it; *nv = sv; return;
    default:         *ns = S_UNDEF; *nt = IT_UNDEF; *nv = SV_UNDEF; return;
    }
}

void mtdo_vacc_multipliers(const mtdo_params *P, int32_t last, int32_t day, double *imm, double *asym)   /* util.py:117-150 */
{
    *imm = 1.0; *asym = 1.0;
    if (last == MTDO_NONE) return;
    int32_t d = day - last;
    if (d == 0) { *imm = 1.0 - P->vacc_immunity; *asym = 1.0 - P->vacc_asymptomatic; return; }
    double x = (double)d / P->vacc_decay_interval;
    if (x < 1) x = 1;
    *imm = 1.0 - pow(P->vacc_immunity, x);
    *asym = 1.0 - pow(P->vacc_asymptomatic, x);
}

/* ------------------------------------------------------------------ replay-tape lookups (binary search on sorted keys) */
static int64_t tape_find_sel(const mtdo_tape *T, int32_t cell, int32_t agent)
{
    uint64_t key = ((uint64_t)(uint32_t)cell << 32) | (uint32_t)agent;
    int64_t lo = 0, hi = T->n_sel;
    while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (T->sel_key[mid] < key) lo = mid + 1; else hi = mid; }
    return (lo < T->n_sel && T->sel_key[lo] == key) ? lo : -1;
}

static int64_t tape_find_pair(const mtdo_tape *T, int32_t cell, int32_t a1, int32_t a2)
{
    uint64_t k1 = ((uint64_t)(uint32_t)cell << 32) | (uint32_t)a1;
    uint32_t k2 = (uint32_t)a2;
    int64_t lo = 0, hi = T->n_pair;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if (T->pair_k1[mid] < k1 || (T->pair_k1[mid] == k1 && T->pair_k2[mid] < k2)) lo = mid + 1; else hi = mid;
    }
    return (lo < T->n_pair && T->pair_k1[lo] == k1 && T->pair_k2[lo] == k2) ? lo : -1;
}

static int64_t tape_find_expo(const mtdo_tape *T, int32_t agent)
{
    int64_t lo = 0, hi = T->n_expo;
    while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (T->expo_agent[mid] < agent) lo = mid + 1; else hi = mid; }
    return (lo < T->n_expo && T->expo_agent[lo] == agent) ? lo : -1;
}

/* ------------------------------------------------------------------ grouping of the emission stream by cell
 * vars_util.cells_agents_timesteps is a dict {cellid: [(agent, start, end), ...]} filled in emission order
 * (itinerary.py:1446-1452): cell order = order of first appearance, list order = emission order. */
typedef struct {
    int64_t n_slots;
    int32_t max_cell;
    int32_t *slot_of_cell;   /* [max_cell+1], -1 = inactive */
    int32_t *cell_of_slot;   /* [n_slots] */
    int64_t *off;            /* [n_slots+1] */
    int64_t *idx;            /* [M] emission indices grouped by slot, stable */
} cellgroups;

static void group_by_cell(int64_t M, const int32_t *iv_cell, cellgroups *G)
{
    int32_t mx = -1;
    for (int64_t i = 0; i < M; i++) if (iv_cell[i] > mx) mx = iv_cell[i];
    G->max_cell = mx;
    G->slot_of_cell = (int32_t *)malloc(sizeof(int32_t) * (size_t)(mx + 2));
    for (int32_t c = 0; c <= mx; c++) G->slot_of_cell[c] = -1;
    int64_t ns = 0;
    for (int64_t i = 0; i < M; i++) if (G->slot_of_cell[iv_cell[i]] < 0) G->slot_of_cell[iv_cell[i]] = (int32_t)ns++;
    G->n_slots = ns;
    G->cell_of_slot = (int32_t *)malloc(sizeof(int32_t) * (size_t)(ns + 1));
    G->off = (int64_t *)calloc((size_t)(ns + 2), sizeof(int64_t));
    G->idx = (int64_t *)malloc(sizeof(int64_t) * (size_t)(M + 1));
    for (int64_t i = 0; i < M; i++) { int32_t s = G->slot_of_cell[iv_cell[i]]; G->cell_of_slot[s] = iv_cell[i]; G->off[s + 1]++; }
    for (int64_t s = 0; s < ns; s++) G->off[s + 1] += G->off[s];
    int64_t *cur = (int64_t *)malloc(sizeof(int64_t) * (size_t)(ns + 1));
    memcpy(cur, G->off, sizeof(int64_t) * (size_t)ns);
    for (int64_t i = 0; i < M; i++) G->idx[cur[G->slot_of_cell[iv_cell[i]]]++] = i;
    free(cur);
}

static void free_groups(cellgroups *G) { free(G->slot_of_cell); free(G->cell_of_slot); free(G->off); free(G->idx); }

/* ------------------------------------------------------------------ generate_contact_network (contactnetwork.py:134-315) */
typedef struct { int32_t a, b; int32_t ov_head, ov_tail; } potpair;    /* key (a,b) in local agent indices */
typedef struct { uint8_t s, e; int32_t next; } ovent;

typedef struct {
    /* per-cell scratch, grown on demand */
    int64_t cap_m, cap_n, cap_nn, cap_pairs, cap_ov, cap_adj;
    int32_t *la;        /* [m] local agent index of interval */
    int32_t *ids;       /* [n] agents_ids (first appearance order) */
    int64_t *ts;        /* [n] agents_total_timesteps */
    int32_t *pc;        /* [n] agents_potentialcontacts_count */
    uint8_t *has_pc;    /* [n] entry exists in agents_potentialcontacts_count */
    double *degf;       /* [n] float degrees */
    int64_t *deg;       /* [n] integer degrees */
    uint8_t *dead;      /* [n] all pairs of this agent were deleted (delete_all_pairs_by_id) */
    int32_t *order;     /* [n] greedy order */
    int32_t *pidx;      /* [n*n] pair index for key (a,b), -1 none */
    potpair *pairs; int64_t n_pairs;
    ovent *ov; int64_t n_ov;
    int32_t *adj_head, *adj_tail;   /* [n] per-agent list of pair indices, insertion order */
    int32_t *adj_next_a, *adj_next_b; /* [pairs] next pair in a's / b's list */
    int32_t *cand;      /* [n] candidate partners */
    int32_t *local_of_agent; int64_t n_agents;
} cellws;

static void ws_reserve(cellws *W, int64_t m, int64_t n)
{
    if (m > W->cap_m) { W->cap_m = m * 2; W->la = (int32_t *)realloc(W->la, sizeof(int32_t) * (size_t)W->cap_m); }
    if (n > W->cap_n) {
        W->cap_n = n * 2;
        size_t c = (size_t)W->cap_n;
        W->ids = (int32_t *)realloc(W->ids, 4 * c); W->ts = (int64_t *)realloc(W->ts, 8 * c);
        W->pc = (int32_t *)realloc(W->pc, 4 * c); W->has_pc = (uint8_t *)realloc(W->has_pc, c);
        W->degf = (double *)realloc(W->degf, 8 * c); W->deg = (int64_t *)realloc(W->deg, 8 * c);
        W->dead = (uint8_t *)realloc(W->dead, c); W->order = (int32_t *)realloc(W->order, 4 * c);
        W->adj_head = (int32_t *)realloc(W->adj_head, 4 * c); W->adj_tail = (int32_t *)realloc(W->adj_tail, 4 * c);
        W->cand = (int32_t *)realloc(W->cand, 4 * c);
    }
    if (n * n > W->cap_nn) { W->cap_nn = n * n * 2; W->pidx = (int32_t *)realloc(W->pidx, 4 * (size_t)W->cap_nn); }
}

static int32_t ws_new_pair(cellws *W, int32_t a, int32_t b)
{
    if (W->n_pairs >= W->cap_pairs) {
        W->cap_pairs = W->cap_pairs ? W->cap_pairs * 2 : 1024;
        W->pairs = (potpair *)realloc(W->pairs, sizeof(potpair) * (size_t)W->cap_pairs);
        W->adj_next_a = (int32_t *)realloc(W->adj_next_a, 4 * (size_t)W->cap_pairs);
        W->adj_next_b = (int32_t *)realloc(W->adj_next_b, 4 * (size_t)W->cap_pairs);
    }
    int32_t p = (int32_t)W->n_pairs++;
    W->pairs[p].a = a; W->pairs[p].b = b; W->pairs[p].ov_head = W->pairs[p].ov_tail = -1;
    W->adj_next_a[p] = W->adj_next_b[p] = -1;
    /* append p to a's and b's per-agent lists (dict insertion order restricted to keys containing the agent) */
    int32_t ends[2] = { a, b };
    for (int k = 0; k < 2; k++) {
        int32_t x = ends[k];
        if (W->adj_head[x] < 0) W->adj_head[x] = p;
        else {
            int32_t t = W->adj_tail[x];
            if (W->pairs[t].a == x) W->adj_next_a[t] = p; else W->adj_next_b[t] = p;
        }
        W->adj_tail[x] = p;
    }
    return p;
}

static void ws_add_overlap(cellws *W, int32_t p, int lo, int hi)
{
    if (W->n_ov >= W->cap_ov) {
        W->cap_ov = W->cap_ov ? W->cap_ov * 2 : 4096;
        W->ov = (ovent *)realloc(W->ov, sizeof(ovent) * (size_t)W->cap_ov);
    }
    int32_t o = (int32_t)W->n_ov++;
    W->ov[o].s = (uint8_t)lo; W->ov[o].e = (uint8_t)hi; W->ov[o].next = -1;
    if (W->pairs[p].ov_head < 0) W->pairs[p].ov_head = o; else W->ov[W->pairs[p].ov_tail].next = o;
    W->pairs[p].ov_tail = o;
}

static int cmp_desc_deg_id(const void *x, const void *y, void *arg)
{
    cellws *W = (cellws *)arg;
    int32_t i = *(const int32_t *)x, j = *(const int32_t *)y;
    if (W->deg[i] != W->deg[j]) return W->deg[i] > W->deg[j] ? -1 : 1;     /* sorted(zip(deg, id), reverse=True) */
    return W->ids[i] > W->ids[j] ? -1 : (W->ids[i] < W->ids[j] ? 1 : 0);
}

/* Test knob (tests/test_device_log.py): the two per-cell logarithms log(avg_ts), log(avg_pc) moved by this many ulps.  The
 * CUDA path computes them with CUDA's log (<= 1 ulp), the reference with glibc's: a degree can only differ if it sits
 * within such a perturbation of a rounding boundary, which the test shows does not happen on any recorded cell. */
int mtdo_log_ulp_shift = 0;
static double log_avg(double x)
{
    double l = log(x);
    for (int i = 0; i < mtdo_log_ulp_shift; i++) l = nextafter(l, INFINITY);
    for (int i = 0; i > mtdo_log_ulp_shift; i--) l = nextafter(l, -INFINITY);
    return l;
}
void mtdo_log_array(int64_t n, const double *x, double *out) { for (int64_t i = 0; i < n; i++) out[i] = log(x[i]); }

/* One cell.  iv = emission indices of the cell's intervals in list order.  Appends rows; returns number appended,
 * or -1 on output overflow. */
static int64_t run_cell(const mtdo_params *P, const mtdo_dayctx *D, cellws *W, int32_t cellid, int64_t m,
                        const int64_t *iv, const int32_t *iv_agent, const uint8_t *iv_start, const uint8_t *iv_end,
                        int ctype, int indid, const mtdo_agents *A, int64_t cap_left, int32_t *rows, int32_t *row_dur,
                        int32_t *row_cell, mtdo_stats *S)
{
    /* agents_ids in order of first appearance (contactnetwork.py:161-163) */
    int64_t n = 0;
    ws_reserve(W, m, m);
    for (int64_t i = 0; i < m; i++) {
        int32_t ag = iv_agent[iv[i]];
        if (W->local_of_agent[ag] < 0) { W->local_of_agent[ag] = (int32_t)n; W->ids[n] = ag; n++; }
        W->la[i] = W->local_of_agent[ag];
    }
    for (int64_t a = 0; a < n; a++) {
        W->ts[a] = 0; W->pc[a] = 0; W->has_pc[a] = 0; W->degf[a] = 0.0; W->deg[a] = 0; W->dead[a] = 0;
        W->adj_head[a] = W->adj_tail[a] = -1;
    }
    for (int64_t q = 0; q < n * n; q++) W->pidx[q] = -1;
    W->n_pairs = 0; W->n_ov = 0;

    int act = mtdo_activity_of_celltype(ctype, ctype == CT_WORKPLACE ? indid : -1);   /* :150-156 */
    int64_t ts_sum = 0, same_agent_sq = 0;

    /* ---- pair enumeration (:160-196) ---- */
    for (int64_t i = 0; i < m; i++) {
        int32_t a = W->la[i];
        int si = iv_start[iv[i]], ei = iv_end[iv[i]];
        int len = ei - si + 1;            /* len(range(start, end+1)) */
        if (len < 0) len = 0;
        ts_sum += len; W->ts[a] += len;
        for (int64_t j = 0; j < m; j++) {
            int32_t b = W->la[j];
            if (b == a) continue;
            if (W->pidx[(int64_t)b * n + a] >= 0) continue;            /* pair_already_computed: (ag_id, agentid) in dict */
            int sj = iv_start[iv[j]], ej = iv_end[iv[j]];
            int lo = si > sj ? si : sj, hi = ei < ej ? ei : ej;        /* get_overlapping_range (:426-432) */
            W->has_pc[a] = 1; W->has_pc[b] = 1;
            if (lo <= hi) {
                int32_t p = W->pidx[(int64_t)a * n + b];
                if (p < 0) {                                            /* pair_exists is false in both orientations */
                    W->pc[a]++; W->pc[b]++;
                    p = ws_new_pair(W, a, b);
                    W->pidx[(int64_t)a * n + b] = p;
                }
                ws_add_overlap(W, p, lo, hi);
            }
        }
    }
    {   /* metric: unordered interval pairs of different agents */
        int64_t *cnt = (int64_t *)calloc((size_t)n, sizeof(int64_t));
        for (int64_t i = 0; i < m; i++) cnt[W->la[i]]++;
        for (int64_t a = 0; a < n; a++) same_agent_sq += cnt[a] * cnt[a];
        free(cnt);
        S->pairs_evaluated += (m * m - same_agent_sq) / 2;
        S->n_potential += W->n_pairs;
    }
    for (int64_t a = 0; a < n; a++) W->local_of_agent[W->ids[a]] = -1;   /* reset the global map */

    int64_t n_rows = 0;
    if (W->n_pairs == 0) return 0;                                      /* :198 */

    /* ---- degrees (:210-269) ---- */
    int64_t k = P->k_by_act[act];
    int64_t max_degree = k * n;                                          /* round(k * n) */
    double avg_ts = (double)ts_sum / (double)n;
    int64_t pc_sum = 0, pc_cnt = 0;
    for (int64_t a = 0; a < n; a++) if (W->has_pc[a]) { pc_sum += W->pc[a]; pc_cnt++; }
    double avg_pc = (double)pc_sum / (double)pc_cnt;                     /* np.mean */
    double one_minus_mhd = 1 - D->mhd;
    for (int64_t a = 0; a < n; a++) {
        if (W->pc[a] > 0) {
            int ab = A->age_bracket[W->ids[a]];
            if (ab < 0) ab += 16;                                        /* numpy negative index */
            double cmv = P->cm[ab][act];
            double tsm = 1.0, pcm = 1.0;
            if ((double)W->ts[a] != avg_ts && avg_ts > 1) tsm = log((double)W->ts[a]) / log_avg(avg_ts);   /* math.log(x, base) */
            if ((double)W->pc[a] != avg_pc && avg_pc > 1) pcm = log((double)W->pc[a]) / log_avg(avg_pc);
            double soc = A->soc_rate[W->ids[a]];
            W->degf[a] = cmv * tsm * pcm * soc * one_minus_mhd;
        }
    }
    int64_t deg_sum = py_round(mtdo_np_pairwise_sum(W->degf, n));
    for (int64_t a = 0; a < n; a++) {
        int64_t d;
        if (deg_sum > max_degree) d = py_round((W->degf[a] / (double)deg_sum) * (double)max_degree);
        else d = py_round(W->degf[a]);
        if (d > W->pc[a]) d = W->pc[a];
        W->deg[a] = d;
    }

    /* ---- greedy sampling (:283-301) ---- */
    for (int64_t a = 0; a < n; a++) W->order[a] = (int32_t)a;
    qsort_r(W->order, (size_t)n, sizeof(int32_t), cmp_desc_deg_id, W);
    for (int64_t oi = 0; oi < n; oi++) {
        int32_t x = W->order[oi];
        int64_t degree = W->deg[x];
        if (degree <= 0) continue;
        /* util.get_all_contacts_ids_by_id: keys in insertion order containing x whose partner has degree > 0 */
        int64_t nc = 0;
        for (int32_t p = W->adj_head[x]; p >= 0; p = (W->pairs[p].a == x) ? W->adj_next_a[p] : W->adj_next_b[p]) {
            int32_t y = (W->pairs[p].a == x) ? W->pairs[p].b : W->pairs[p].a;
            if (W->dead[x] || W->dead[y]) continue;                     /* key was deleted */
            if (W->deg[y] > 0) W->cand[nc++] = y;
        }
        if (nc == 0) continue;                                          /* returns None (:294) */
        if (degree > nc) degree = nc;
        /* np.random.shuffle + np.random.choice(size=degree, replace=False): replay or Philox */
        int32_t picks_local[1];
        (void)picks_local;
        const int32_t *tp = NULL; int64_t ntp = 0;
        if (D->mode == MTDO_MODE_REPLAY) {
            int64_t r = tape_find_sel(D->tape, cellid, W->ids[x]);
            if (r < 0) { S->tape_miss++; continue; }
            tp = D->tape->sel_picks + D->tape->sel_off[r];
            ntp = D->tape->sel_off[r + 1] - D->tape->sel_off[r];
            if (ntp != degree) { S->tape_miss++; if (ntp > degree) ntp = degree; }
        } else {
            /* Philox mode (this project's own draw scheme; the distribution equals shuffle + choice(replace=False):
             * a uniformly random ordered `degree`-subset).  The candidates are put in CANONICAL order (ascending
             * local agent index = order of first appearance in the cell) and draw t takes the r_t-th REMAINING
             * candidate, r_t = floor(u_t * (nc - t)); the taken candidate is removed, order kept. */
            for (int64_t i = 1; i < nc; i++) {                           /* insertion sort by local index */
                int32_t v = W->cand[i]; int64_t j = i - 1;
                while (j >= 0 && W->cand[j] > v) { W->cand[j + 1] = W->cand[j]; j--; }
                W->cand[j + 1] = v;
            }
            for (int64_t t = 0; t < degree; t++) {
                uint32_t w[4];
                philox_site(D, MTDO_SITE_SELECT, (uint32_t)cellid, (uint32_t)W->ids[x], (uint32_t)(t >> 2), w);
                int64_t r = t + (int64_t)(((uint64_t)w[t & 3] * (uint64_t)(nc - t)) >> 32);
                int32_t y = W->cand[r];                                  /* cand[t..nc) = remaining, ascending */
                for (int64_t j = r; j > t; j--) W->cand[j] = W->cand[j - 1];
                W->cand[t] = y;
            }
            ntp = degree;
        }
        /* create_contacts (:386-424) */
        for (int64_t t = 0; t < ntp; t++) {
            int32_t y;
            if (tp) { int32_t pi = tp[t]; if (pi < 0 || pi >= nc) { S->tape_miss++; continue; } y = W->cand[pi]; }
            else y = W->cand[t];
            int32_t p = W->pidx[(int64_t)x * n + y];
            if (p < 0) p = W->pidx[(int64_t)y * n + x];
            if (n_rows >= cap_left) return -1;
            int32_t *r = rows + 5 * n_rows;
            int dur = 0;
            for (int32_t o = W->pairs[p].ov_head; o >= 0; o = W->ov[o].next) dur += W->ov[o].e - W->ov[o].s + 1;
            r[0] = mtdo_simcelltype(ctype);
            r[1] = W->ids[W->pairs[p].a]; r[2] = W->ids[W->pairs[p].b];
            r[3] = W->ov[W->pairs[p].ov_head].s;                        /* first element's start (:101, quirk B7) */
            r[4] = W->ov[W->pairs[p].ov_tail].e;                        /* last element's end */
            row_dur[n_rows] = dur; row_cell[n_rows] = cellid;
            n_rows++;
            W->deg[y] -= 1;
            if (W->deg[y] == 0) W->dead[y] = 1;                          /* delete_all_pairs_by_id(contact) */
        }
        W->deg[x] = degree - ntp;                                        /* main_agent_degree -= len(direct_contacts_ids) */
        if (W->deg[x] == 0) W->dead[x] = 1;
    }
    return n_rows;
}

int64_t mtdo_contact_network(const mtdo_params *P, const mtdo_dayctx *D, int64_t M, const int32_t *iv_cell,
                             const int32_t *iv_agent, const uint8_t *iv_start, const uint8_t *iv_end,
                             const int8_t *cell_type, const int16_t *cell_indid, const mtdo_agents *A,
                             int64_t cap, int32_t *rows, int32_t *row_dur, int32_t *row_cell, mtdo_stats *S)
{
    cellgroups G;
    group_by_cell(M, iv_cell, &G);
    cellws W;
    memset(&W, 0, sizeof(W));
    W.n_agents = A->n_agents;
    W.local_of_agent = (int32_t *)malloc(sizeof(int32_t) * (size_t)(A->n_agents + 1));
    for (int64_t i = 0; i < A->n_agents; i++) W.local_of_agent[i] = -1;
    int64_t total = 0;
    S->n_cells += G.n_slots;
    for (int64_t s = 0; s < G.n_slots; s++) {
        int32_t cellid = G.cell_of_slot[s];
        int64_t m = G.off[s + 1] - G.off[s];
        int64_t r = run_cell(P, D, &W, cellid, m, G.idx + G.off[s], iv_agent, iv_start, iv_end, cell_type[cellid],
                             cell_indid ? cell_indid[cellid] : -1, A, cap - total, rows + 5 * total, row_dur + total,
                             row_cell + total, S);
        if (r < 0) { total = -1; break; }
        total += r;
    }
    free(W.la); free(W.ids); free(W.ts); free(W.pc); free(W.has_pc); free(W.degf); free(W.deg); free(W.dead);
    free(W.order); free(W.pidx); free(W.pairs); free(W.ov); free(W.adj_head); free(W.adj_tail); free(W.adj_next_a);
    free(W.adj_next_b); free(W.cand); free(W.local_of_agent);
    free_groups(&G);
    return total;
}

/* ------------------------------------------------------------------ draws of one exposure (replay row or Philox) */
typedef struct { const mtdo_dayctx *D; int32_t agent; const double *row; mtdo_stats *S; } expodraws;

static double ed_uniform(expodraws *E, int slot)
{
    if (E->D->mode == MTDO_MODE_REPLAY) {
        if (!E->row || isnan(E->row[slot])) { E->S->tape_miss++; return 1.0; }
        return E->row[slot];
    }
    uint32_t w[4];
    philox_site(E->D, MTDO_SITE_EXPO, (uint32_t)E->agent, (uint32_t)slot, 0, w);
    return mtdo_u53(w[0], w[1]);
}

/* util.sample_log_normal(mean, std, 1, isInt=True): np.random.lognormal then Python round (util.py:88-115) */
static int32_t ed_lognormal_days(expodraws *E, int slot, double mu, double sigma)
{
    if (E->D->mode == MTDO_MODE_REPLAY) {
        if (!E->row || isnan(E->row[slot])) { E->S->tape_miss++; return 0; }
        return (int32_t)E->row[slot];
    }
    uint32_t w[4];
    philox_site(E->D, MTDO_SITE_EXPO, (uint32_t)E->agent, (uint32_t)slot, 0, w);
    double u1 = mtdo_u53(w[0], w[1]), u2 = mtdo_u53(w[2], w[3]);
    double z = sqrt(-2.0 * log(1.0 - u1)) * cos(6.283185307179586 * u2);      /* Box-Muller */
    return (int32_t)py_round(exp(mu + sigma * z));
}

static int32_t ed_timestep_option(expodraws *E, int slot)   /* np.random.choice(np.arange(42, 121)) (epidemiology.py:84) */
{
    if (E->D->mode == MTDO_MODE_REPLAY) {
        if (!E->row || isnan(E->row[slot])) { E->S->tape_miss++; return 42; }
        return (int32_t)E->row[slot];
    }
    uint32_t w[4];
    philox_site(E->D, MTDO_SITE_EXPO, (uint32_t)E->agent, (uint32_t)slot, 0, w);
    return 42 + (int32_t)(((uint64_t)w[0] * 79u) >> 32);
}

static void sched_append(mtdo_agents *A, int32_t ag, int32_t day, int trans, int ts, mtdo_stats *S)
{
    uint8_t c = A->sched_cnt[ag];
    if (c >= MTDO_MAXSCHED) { S->sched_overflow++; return; }
    A->sched[(int64_t)ag * MTDO_MAXSCHED + c] = (int32_t)(((uint32_t)day << 16) | ((uint32_t)trans << 8) | (uint32_t)ts);
    A->sched_cnt[ag] = c + 1;
}

/* incr0: the exposure day (today for a contact of the day step; a day before the start of the simulation for the
 * itinerary's initial cases, itinerary.py:373-395 -> reset_states, epidemiology.py:243); row: the replay draws of this
 * exposure (NULL: looked up in the day's tape); out8 (optional): [seir, infection type, severity, recovered, entries
 * appended, exposure timestep, exposure day, symptomatic day] */
static void expose_agent_ex(const mtdo_params *P, const mtdo_dayctx *D, mtdo_agents *A, int32_t ag, int n_ts,
                            const uint8_t *ts_list, double asym_mult, mtdo_stats *S, int32_t incr0, const double *row,
                            int32_t *out8)
{
    expodraws E = { D, ag, row, S };
    if (D->mode == MTDO_MODE_REPLAY && !row) {
        int64_t r = tape_find_expo(D->tape, ag);
        if (r >= 0) E.row = D->tape->expo_row + r * MTDO_EXPO_W; else S->tape_miss++;
    }
    int32_t current_day = D->day, incr = incr0;
    int first_T = -1, n_app = 0;
    int b = A->epi_age_bracket[ag];
    /* sampled_exposed_timestep = np.random.choice(overlapping_timesteps) (:246) */
    int ts;
    if (D->mode == MTDO_MODE_REPLAY) ts = (E.row && !isnan(E.row[0])) ? (int)E.row[0] : 0;
    else {
        uint32_t w[4];
        philox_site(D, MTDO_SITE_EXPO, (uint32_t)ag, 0, 0, w);
        ts = ts_list[(int64_t)(((uint64_t)w[0] * (uint64_t)n_ts) >> 32)];
    }
    int symptomatic_day = -1, recovered = 0, start_hosp = MTDO_NONE, end_hosp = MTDO_NONE;
    int inf_type;
#define STEP(T) do { incr += ed_lognormal_days(&E, 1 + (T), P->ln_mu[T], P->ln_sigma[T]); \
                     if (incr >= current_day) { if (n_app++ == 0) first_T = (T); sched_append(A, ag, incr, (T), ts, S); } } while (0)
    STEP(T_EXP_INF);                                                                   /* :248-256 */
    double u = ed_uniform(&E, 11);                                                    /* :268 */
    if (u < P->probs_by_age[1][b] * asym_mult) {
        inf_type = IT_PRESYMP;
        STEP(T_INF_SYMP); symptomatic_day = incr;                                      /* :275-281 */
        if (ed_uniform(&E, 12) < P->probs_by_age[2][b]) {                              /* :293-297 */
            STEP(T_SYMP_SEV); start_hosp = incr;
            if (ed_uniform(&E, 13) < P->probs_by_age[3][b]) {                          /* :317-321 */
                STEP(T_SEV_CRI);
                if (ed_uniform(&E, 14) < P->probs_by_age[4][b]) { STEP(T_CRI_DEATH); end_hosp = incr; }   /* :339-347 */
                else { STEP(T_CRI_REC); end_hosp = incr; recovered = 1; }
            } else { STEP(T_SEV_REC); end_hosp = incr; recovered = 1; }
        } else { STEP(T_MILD_REC); recovered = 1; }
    } else {
        inf_type = IT_PREASYMP;
        STEP(T_ASYMP_REC); recovered = 1;                                              /* :428-447 */
    }
    if (recovered) STEP(T_REC_SUS);                                                    /* :449-455 */
#undef STEP
    if (start_hosp != MTDO_NONE) {                                                     /* schedule_hospitalisation (:647-660) */
        int32_t *h = A->hosp + (int64_t)ag * 3;
        if (h[0] != MTDO_NONE) { h[1] = ts; h[2] = end_hosp; }                         /* keeps the earlier start day */
        else { h[0] = start_hosp; h[1] = ts; h[2] = end_hosp; }
    }
    if (symptomatic_day > -1) {
        int tts = ed_timestep_option(&E, 15);                                          /* :480 */
        if (D->testing_enabled) {                                                      /* schedule_test (:502-565), Symptomatic */
            int32_t *td = A->test_day + (int64_t)ag * 2, *tr = A->test_result_day + (int64_t)ag * 2;
            if (td[0] == MTDO_NONE) {
                if (ed_uniform(&E, 16) < P->testing_after_symptoms_probability) {
                    int32_t testing_day = symptomatic_day + ed_lognormal_days(&E, 17, P->test_mu, P->test_sigma);
                    td[0] = testing_day; td[1] = tts;
                    tr[0] = testing_day + ed_lognormal_days(&E, 18, P->result_mu, P->result_sigma); tr[1] = tts;
                }
            }
        }
        (void)ed_timestep_option(&E, 19);                                              /* :496 */
        if (D->quarantine_enabled) {
            /* schedule_quarantine(Symptomatic) consumes one uniform (:575); whatever it stores in
             * agent["quarantine_days"] is overwritten by simulate_direct_contacts with the value read before the
             * call (epidemiology.py:206,216) -> no state change. */
            (void)ed_uniform(&E, 20);
        }
    }
    int seir = S_EXP, it = inf_type, sv = SV_UNDEF;                                    /* :219-225, :238-240 */
    if (incr0 < 1) {
        /* reset_states (:243): the states that PRECEDE the first transition still ahead (:257-470, as written: while
         * waiting for ExposedToInfectious the type is PreAsymptomatic whatever the branch, :262-263; nothing ahead ->
         * the agent stays Exposed with an empty schedule) */
        switch (first_T) {
        case T_EXP_INF:  seir = S_EXP; it = IT_PREASYMP; sv = SV_UNDEF; break;
        case T_INF_SYMP: seir = S_INF; it = IT_PRESYMP; sv = SV_UNDEF; break;
        case T_SYMP_SEV: case T_MILD_REC: seir = S_INF; it = IT_SYMP; sv = SV_MILD; break;
        case T_SEV_CRI:  case T_SEV_REC:  seir = S_INF; it = IT_SYMP; sv = SV_SEVERE; break;
        case T_CRI_DEATH: case T_CRI_REC: seir = S_INF; it = IT_SYMP; sv = SV_CRITICAL; break;
        case T_ASYMP_REC: seir = S_INF; it = IT_ASYMP; sv = SV_UNDEF; break;
        case T_REC_SUS:  seir = S_REC; it = IT_UNDEF; sv = SV_UNDEF; break;
        default: break;
        }
    }
    A->seir[ag] = (int8_t)seir; A->inf_type[ag] = (int8_t)it; A->inf_sev[ag] = (int8_t)sv;
    if (out8) {
        out8[0] = seir; out8[1] = it; out8[2] = sv; out8[3] = recovered; out8[4] = n_app; out8[5] = ts; out8[6] = incr0;
        out8[7] = symptomatic_day;
    }
}

/* simulate_seir_state_transition + schedule_* as reached from simulate_direct_contacts (epidemiology.py:233-660) */
static void expose_agent(const mtdo_params *P, const mtdo_dayctx *D, mtdo_agents *A, int32_t ag, int n_ts,
                         const uint8_t *ts_list, double asym_mult, mtdo_stats *S)
{
    expose_agent_ex(P, D, A, ag, n_ts, ts_list, asym_mult, S, D->day, NULL, NULL);
}

/* the itinerary stage's initial cases (itinerary.py:373-395): Epidemiology.simulate_seir_state_transition called with an
 * exposure day before day 1.  exposure_day NULL: drawn from [day - pre_days, day) with word 1 of the slot-0 Philox
 * block (Philox mode).  draws: replay rows, n x MTDO_EXPO_W.  Returns 0, or -1 on a bad id. */
int64_t mtdo_seed_exposures(const mtdo_params *P, const mtdo_dayctx *D, mtdo_agents *A, int64_t n, const int32_t *ids,
                            const int32_t *exposure_day, int32_t pre_days, const uint8_t *ts_list, int32_t n_ts,
                            const double *asym_mult, const double *draws, int32_t *out8, mtdo_stats *S)
{
    for (int64_t i = 0; i < n; i++) {
        const int32_t ag = ids[i];
        if (ag < 0 || ag >= A->n_agents) return -1;
        int32_t incr0;
        if (exposure_day) incr0 = exposure_day[i];
        else {
            uint32_t w[4];
            philox_site(D, MTDO_SITE_EXPO, (uint32_t)ag, 0, 0, w);
            incr0 = D->day - pre_days + (int32_t)(((uint64_t)w[1] * (uint64_t)(uint32_t)pre_days) >> 32);
        }
        expose_agent_ex(P, D, A, ag, n_ts, ts_list, asym_mult ? asym_mult[i] : 1.0, S, incr0,
                        draws ? draws + i * MTDO_EXPO_W : NULL, out8 ? out8 + 8 * i : NULL);
    }
    return 0;
}

int64_t mtdo_transmission(const mtdo_params *P, const mtdo_dayctx *D, int64_t M, const int32_t *iv_cell,
                          const int32_t *iv_agent, const uint8_t *iv_start, const uint8_t *iv_end,
                          const int8_t *cell_type, int64_t n_rows, const int32_t *rows, const int32_t *row_dur,
                          const int32_t *row_cell, mtdo_agents *A, int64_t cap_e, int32_t *exposed, double *p_out,
                          mtdo_stats *S)
{
    cellgroups G;
    group_by_cell(M, iv_cell, &G);
    int64_t n_exp = 0;
    uint8_t *tsbuf = (uint8_t *)malloc(144 * 64);
    int64_t tscap = 144 * 64;
    for (int64_t r = 0; r < n_rows; r++) {
        if (p_out) p_out[r] = NAN;
        int32_t a1 = rows[5 * r + 1], a2 = rows[5 * r + 2], cellid = row_cell[r];
        int s1 = A->seir[a1], s2 = A->seir[a2];
        if (!((s1 == S_INF && s2 == S_SUS) || (s1 == S_SUS && s2 == S_INF))) continue;      /* :106-107 */
        int dur = row_dur[r];
        if (dur <= 0) continue;
        int32_t ex = (s1 == S_SUS) ? a1 : a2;
        double susc = P->probs_by_age[0][A->epi_age_bracket[ex]];
        int sct = mtdo_simcelltype(cell_type[cellid]);
        double base;                                                                     /* :998-1014 */
        if (sct == SCT_RESIDENCE) base = P->p_household;
        else if (sct == SCT_WORKPLACE) base = (cellid == A->res_cellid[ex]) ? P->p_workplace : P->p_community;  /* quirk B4 */
        else if (sct == SCT_SCHOOL) base = P->p_school;
        else base = P->p_community;
        double lg = log((double)dur);
        double mult = (lg > 1.0) ? lg : 1.0;                                             /* max(1, math.log(duration)) */
        double imm, asym;
        mtdo_vacc_multipliers(P, A->last_dose_day[ex], D->day, &imm, &asym);
        double p = base * (mult * susc * (1 - D->mhd) * imm);                            /* :196 */
        if (p_out) p_out[r] = p;
        double u;
        if (D->mode == MTDO_MODE_REPLAY) {
            int64_t t = tape_find_pair(D->tape, cellid, a1, a2);
            if (t < 0) { S->tape_miss++; continue; }
            u = D->tape->pair_u[t];
        } else {
            uint32_t w[4];
            philox_site(D, MTDO_SITE_PAIR, (uint32_t)cellid, (uint32_t)a1, (uint32_t)a2, w);
            u = mtdo_u53(w[0], w[1]);
        }
        if (!(u < p)) continue;
        /* overlapping timesteps, flattened in overlap-list order: a1's intervals (outer) x a2's (inner) */
        int64_t nts = 0;
        int32_t slot = G.slot_of_cell[cellid];
        for (int64_t i = G.off[slot]; i < G.off[slot + 1]; i++) {
            if (iv_agent[G.idx[i]] != a1) continue;
            for (int64_t j = G.off[slot]; j < G.off[slot + 1]; j++) {
                if (iv_agent[G.idx[j]] != a2) continue;
                int lo = iv_start[G.idx[i]] > iv_start[G.idx[j]] ? iv_start[G.idx[i]] : iv_start[G.idx[j]];
                int hi = iv_end[G.idx[i]] < iv_end[G.idx[j]] ? iv_end[G.idx[i]] : iv_end[G.idx[j]];
                for (int t = lo; t <= hi; t++) {
                    if (nts >= tscap) { tscap *= 2; tsbuf = (uint8_t *)realloc(tsbuf, (size_t)tscap); }
                    tsbuf[nts++] = (uint8_t)t;
                }
            }
        }
        if (n_exp >= cap_e) { n_exp = -1; break; }
        exposed[n_exp++] = ex;
        expose_agent(P, D, A, ex, (int)nts, tsbuf, asym, S);
    }
    free(tsbuf);
    free_groups(&G);
    return n_exp;
}

/* ------------------------------------------------------------------ seirstateutil.update_agent_state (seirstateutil.py:38-72) */
/* Interval builder: Itinerary.update_cell_agents_timesteps + combine_same_cell_itinerary_entries
 * (itinerary.py:1403-1488).  Itineraries as CSR (entries sorted by timestep; cell -1 = residence) + agent groups.
 * Writes the interval stream in emission order; returns its length (or -1 if cap is too small). */
int64_t mtdo_build_intervals(int64_t n_itin, const int64_t *ent_off, const int16_t *ent_ts, const int32_t *ent_cell,
                             const int64_t *grp_off, const int32_t *grp_agent, const int32_t *grp_rescell, int64_t cap,
                             int32_t *iv_agent, int32_t *iv_cell, uint8_t *iv_start, uint8_t *iv_end)
{
    int64_t out = 0;
    for (int64_t t = 0; t < n_itin; t++) {
        /* combine_same_cell_itinerary_entries (:1459-1488): keep the first entry of every run of equal cells */
        int64_t kept[512]; int nk = 0;
        for (int64_t k = ent_off[t]; k < ent_off[t + 1]; k++)
            if (nk == 0 || ent_cell[k] != ent_cell[kept[nk - 1]]) { if (nk < 512) kept[nk++] = k; }
        for (int i = 0; i < nk; i++) {
            int start = ent_ts[kept[i]], end;                                /* :1413-1433 */
            if (i == 0) { start = 0; end = 143; if (nk > 1) end = ent_ts[kept[1]]; }
            else if (i == nk - 1) end = 143;
            else end = ent_ts[kept[i + 1]];
            if (start == end) continue;                                      /* :1435 */
            int32_t cell = ent_cell[kept[i]];
            for (int64_t g = grp_off[t]; g < grp_off[t + 1]; g++) {
                if (cell == -1) cell = grp_rescell[g];                       /* :1439-1440: sticks after the first agent */
                if (out >= cap) return -1;
                iv_agent[out] = grp_agent[g]; iv_cell[out] = cell;
                iv_start[out] = (uint8_t)(start < 0 ? 0 : start); iv_end[out] = (uint8_t)(end > 143 ? 143 : (end < 0 ? 0 : end));
                out++;
            }
        }
    }
    return out;
}

/* Vaccination scheduling (Epidemiology.schedule_vaccinations, epidemiology.py:911-976), device part: k locals without a
 * dose, drawn without replacement, get the record [day + 1, timestep].  Philox mode: the k smallest keys
 * (philox(day, agent; site 7).w0 << 32 | agent) -- a uniform k-subset like np.random.choice(replace=False); timestep =
 * 42 + floor(w1 * 79 / 2^32).  selected[a] = 1 for the agents scheduled.  Returns their number. */
static int cmp_u64(const void *a, const void *b)
{
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}

int64_t mtdo_schedule_vaccinations(const mtdo_dayctx *D, int64_t n_locals, int64_t k, const int32_t *last_dose_day,
                                   int32_t *vacc_day, uint8_t *selected)
{
    uint64_t *keys = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(n_locals > 0 ? n_locals : 1));
    int64_t nc = 0;
    for (int64_t a = 0; a < n_locals; a++) {
        selected[a] = 0;
        if (last_dose_day[a] != MTDO_NONE) continue;
        uint32_t ctr[4] = { (uint32_t)D->day, (uint32_t)a, 0u, 0u }, key[2] = { D->seed, 7u }, w[4];
        mtdo_philox4x32(ctr, key, w);
        keys[nc++] = ((uint64_t)w[0] << 32) | (uint64_t)(uint32_t)a;
    }
    if (k > nc) k = nc;
    if (k < 0) k = 0;
    qsort(keys, (size_t)nc, sizeof(uint64_t), cmp_u64);
    for (int64_t i = 0; i < k; i++) {
        const int64_t a = (int64_t)(keys[i] & 0xffffffffu);
        uint32_t ctr[4] = { (uint32_t)D->day, (uint32_t)a, 0u, 0u }, key[2] = { D->seed, 7u }, w[4];
        mtdo_philox4x32(ctr, key, w);
        vacc_day[2 * a] = D->day + 1;
        vacc_day[2 * a + 1] = 42 + (int32_t)(((uint64_t)w[1] * 79u) >> 32);
        selected[a] = 1;
    }
    free(keys);
    return k;
}

/* dose bookkeeping of the vaccination day (itinerary.py:2489-2527) */
void mtdo_vaccination_day(int32_t day, int64_t n_locals, int32_t *last_dose_day, int32_t *vacc_day)
{
    for (int64_t a = 0; a < n_locals; a++) {
        const int32_t vd = vacc_day[2 * a];
        if (vd == MTDO_NONE) continue;
        if (day == vd) last_dose_day[a] = day;
        else if (day > vd) { vacc_day[2 * a] = MTDO_NONE; vacc_day[2 * a + 1] = MTDO_NONE; }
    }
}

/* Intervention bookkeeping of the itinerary stage, once per agent and day (Itinerary.sample_intervention_activities,
 * itinerary.py:2283-2528), in the reference's order:
 *   (A) :2301-2332  a test result due today flags the agent positive unless a false negative (Exposed / Infectious) or
 *       only on a false positive (everybody else); a positive agent enters contact_tracing_agent_ids and gets a Positive
 *       quarantine (schedule_quarantine, epidemiology.py:568-625); a result day in the past is cleared;
 *   (B) :2334-2397  hospitalisation_days are cleared on their end day (not on a day that is also their start day);
 *   (C) :2399-2457  quarantine_days are cleared on their end day, unless the window starts today or the hospitalisation
 *       ended today inside the window (the agent is then sent home for the rest of it);
 *   (D) :2459-2487  test_day is cleared six days after the test.
 * The stage's re-timing of records on arrival / departure days (:2342-2383, :2412-2425, :2467-2481) belongs to the
 * itinerary and is not part of this path.  Draws: replay (rp_agent ascending, two doubles per record: the result draw and
 * the quarantine draw, NaN = not consumed) or Philox keyed (day, agent), site 6.
 * positive[a] = 1 for the agents added to contact_tracing_agent_ids.  Returns their number. */
int64_t mtdo_process_interventions(const mtdo_params *P, const mtdo_dayctx *D, mtdo_agents *A, const uint8_t *epi_present,
                                   int64_t rp_n, const int32_t *rp_agent, const double *rp_u, uint8_t *positive,
                                   int64_t *tape_miss)
{
    int64_t n = 0;
    for (int64_t a = 0; a < A->n_agents; a++) {
        positive[a] = 0;
        if (epi_present && !epi_present[a]) continue;
        int32_t *tr = A->test_result_day + 2 * a;
        if (tr[0] != MTDO_NONE) {                                       /* (A) */
            if (D->day == tr[0]) {
                const int ts = tr[1];
                double u, u2;
                if (D->mode == MTDO_MODE_REPLAY) {
                    int64_t lo = 0, hi = rp_n;
                    while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (rp_agent[mid] < (int32_t)a) lo = mid + 1; else hi = mid; }
                    if (lo < rp_n && rp_agent[lo] == (int32_t)a) { u = rp_u[2 * lo]; u2 = rp_u[2 * lo + 1]; }
                    else { u = 2.0; u2 = 2.0; if (tape_miss) (*tape_miss)++; }
                } else {
                    uint32_t ctr[4] = { (uint32_t)D->day, (uint32_t)a, 0u, 0u }, key[2] = { D->seed, 6u }, w[4];
                    mtdo_philox4x32(ctr, key, w);
                    u = mtdo_u53(w[0], w[1]); u2 = mtdo_u53(w[2], w[3]);
                }
                int pos;
                if (A->seir[a] == S_EXP || A->seir[a] == S_INF) pos = !(u < P->testing_false_negative_rate);
                else pos = (u < P->testing_false_positive_rate);
                if (pos) {
                    positive[a] = 1; n++;
                    if (D->quarantine_enabled && !(u2 < P->ignore_quarantine_rules_probability)) {
                        int32_t *q = A->quarantine + 3 * a;
                        const int start = D->day, end = D->day + P->quarantine_positive_duration;
                        if (!(q[0] != MTDO_NONE && start >= q[0])) { q[0] = start; q[1] = ts; q[2] = end; }   /* epidemiology.py:606-611 */
                    }
                }
            } else if (D->day > tr[0]) { tr[0] = MTDO_NONE; tr[1] = MTDO_NONE; }
        }
        int hosp_end_today = 0;
        int32_t *hp = A->hosp + 3 * a;
        if (hp[0] != MTDO_NONE) {                                       /* (B) */
            if (hp[0] == D->day) { /* the itinerary sends the agent to a hospital cell */ }
            else if (hp[2] == D->day) { hosp_end_today = 1; hp[0] = hp[1] = hp[2] = MTDO_NONE; }
        }
        int32_t *q = A->quarantine + 3 * a;
        if (q[0] != MTDO_NONE) {                                        /* (C) */
            if (q[0] == D->day || (hosp_end_today && D->day <= q[2])) { /* the itinerary sends the agent home */ }
            else if (q[2] == D->day) { q[0] = q[1] = q[2] = MTDO_NONE; }
        }
        int32_t *td = A->test_day + 2 * a;
        if (td[0] != MTDO_NONE) {                                       /* (D) */
            if (D->day == td[0]) { /* the itinerary adds the visit to a testing hub */ }
            else if (D->day - td[0] == 6) { td[0] = td[1] = MTDO_NONE; }
        }
    }
    return n;
}

int64_t mtdo_process_test_results(const mtdo_params *P, const mtdo_dayctx *D, mtdo_agents *A, const uint8_t *epi_present,
                                  uint8_t *positive)
{
    return mtdo_process_interventions(P, D, A, epi_present, 0, NULL, NULL, positive, NULL);
}

/* Intervention coupling of the tiled configurations: the itinerary stage keeps agents
 * who are deceased, inside a quarantine window or inside a hospitalisation window out of every cell but their residence
 * (itinerary.py:2283-2528 works at timestep granularity; here the whole day is dropped).  keep[i] = 1 if interval i
 * stays in the day's stream.  Returns the number of kept intervals. */
int64_t mtdo_mask_intervals(int32_t day, int64_t M, const int32_t *iv_agent, const int32_t *iv_cell, const mtdo_agents *A,
                            uint8_t *keep)
{
    int64_t n = 0;
    for (int64_t i = 0; i < M; i++) {
        const int32_t a = iv_agent[i];
        const int32_t *q = A->quarantine + 3 * (int64_t)a, *h = A->hosp + 3 * (int64_t)a;
        int masked = A->seir[a] == S_DEC;
        if (q[0] != MTDO_NONE && q[0] <= day && day <= q[2]) masked = 1;
        if (h[0] != MTDO_NONE && h[0] <= day && day <= h[2]) masked = 1;
        keep[i] = !(masked && iv_cell[i] != A->res_cellid[a]);
        n += keep[i];
    }
    return n;
}

int64_t mtdo_advance(int32_t day, mtdo_agents *A, const uint8_t *mask)
{
    int64_t fired = 0;
    for (int64_t ag = 0; ag < A->n_agents; ag++) {
        if (mask && !mask[ag]) continue;
        int c = A->sched_cnt[ag];
        int32_t *sc = A->sched + ag * MTDO_MAXSCHED;
        int today = -1;
        for (int i = 0; i < c; i++) if ((sc[i] >> 16) == day) { today = i; break; }   /* first entry dated today (quirk B3) */
        if (today < 0) continue;
        int trans = (sc[today] >> 8) & 0xff;
        int ns, nt, nv;
        mtdo_transition(A->seir[ag], A->inf_type[ag], A->inf_sev[ag], trans, &ns, &nt, &nv);
        A->seir[ag] = (int8_t)ns; A->inf_type[ag] = (int8_t)nt; A->inf_sev[ag] = (int8_t)nv;
        for (int i = today; i + 1 < c; i++) sc[i] = sc[i + 1];                          /* del list[today_index] */
        A->sched_cnt[ag] = (uint8_t)(c - 1);
        fired++;
    }
    return fired;
}

/* ------------------------------------------------------------------ contact tracing (epidemiology.py:667-834)
 * Draw slots of one primary contact p in Philox mode (site MTDO_SITE_TRACE_PRIM, counter (day, p, slot, 0)):
 *   0 quarantine timestep, 1 quarantine uniform, 2 test timestep, 3 testing days, 4 result days,
 *   16 + t        : t-th draw of the secondary selection (rank among the remaining members),
 *   65536 + 8 j + {0 ts, 1 u, 2 ts, 3 testing days, 4 result days} : j-th selected secondary contact.
 * In replay mode the same draws are consumed positionally from the primary's tape record. */
typedef struct {
    const mtdo_dayctx *D; const mtdo_params *P; int32_t p; const double *rec; int64_t len, pos; mtdo_trace_stats *S;
} tracedraws;

static double td_next(tracedraws *T) { if (T->pos >= T->len) { T->S->tape_miss++; return 0.0; } return T->rec[T->pos++]; }

static int td_ts(tracedraws *T, uint32_t slot)
{
    if (T->D->mode == MTDO_MODE_REPLAY) return (int)td_next(T);
    uint32_t w[4]; philox_site(T->D, MTDO_SITE_TRACE_PRIM, (uint32_t)T->p, slot, 0, w);
    return 42 + (int)(((uint64_t)w[0] * 79u) >> 32);
}
static double td_u(tracedraws *T, uint32_t slot)
{
    if (T->D->mode == MTDO_MODE_REPLAY) return td_next(T);
    uint32_t w[4]; philox_site(T->D, MTDO_SITE_TRACE_PRIM, (uint32_t)T->p, slot, 0, w);
    return mtdo_u53(w[0], w[1]);
}
static int td_ln(tracedraws *T, uint32_t slot, double mu, double sigma)
{
    if (T->D->mode == MTDO_MODE_REPLAY) return (int)td_next(T);
    uint32_t w[4]; philox_site(T->D, MTDO_SITE_TRACE_PRIM, (uint32_t)T->p, slot, 0, w);
    double u1 = mtdo_u53(w[0], w[1]), u2 = mtdo_u53(w[2], w[3]);
    double z = sqrt(-2.0 * log(1.0 - u1)) * cos(6.283185307179586 * u2);
    return (int)py_round(exp(mu + sigma * z));
}

/* schedule_quarantine (epidemiology.py:568-625) for PositiveContact (2) / SecondaryContact (3) on record `r` */
static void trace_quarantine(tracedraws *T, mtdo_agents *A, int32_t r, int day, int ts, int qtype, uint32_t uslot)
{
    if (!T->D->quarantine_enabled) return;
    double u = td_u(T, uslot);
    if (u < T->P->ignore_quarantine_rules_probability) return;
    int start = day, end;
    if (qtype == Q_POSITIVE_CONTACT) { start += T->P->contact_tracing_positive_delay_days; end = start + T->P->quarantine_positive_contact_duration; }
    else { start += T->P->contact_tracing_secondary_delay_days; end = start + T->P->quarantine_secondary_contact_duration; }
    int32_t *q = A->quarantine + 3 * (int64_t)r;
    if (q[0] != MTDO_NONE && start >= q[0]) return;          /* not rescheduled at the same or a later date (:606-611) */
    q[0] = start; q[1] = ts; q[2] = end;
}

/* schedule_test (epidemiology.py:502-565) for PositiveContact / SecondaryContact on record `r` */
static void trace_test(tracedraws *T, mtdo_agents *A, int32_t r, int day, int ts, int qtype, uint32_t lnslot)
{
    if (!T->D->testing_enabled) return;
    int32_t *td = A->test_day + 2 * (int64_t)r, *tr = A->test_result_day + 2 * (int64_t)r;
    if (td[0] != MTDO_NONE) return;                            /* a test is already scheduled (:519-521) */
    int days = (qtype == Q_POSITIVE_CONTACT) ? T->P->contact_tracing_positive_delay_days : T->P->contact_tracing_secondary_delay_days;
    days += td_ln(T, lnslot, T->P->test_mu, T->P->test_sigma);
    int tday = day + days;
    int rday = tday + td_ln(T, lnslot + 1, T->P->result_mu, T->P->result_sigma);
    td[0] = tday; td[1] = ts; tr[0] = rday; tr[1] = ts;
}

static int64_t trace_find_scan(const mtdo_trace_tape *T, uint64_t key)
{
    int64_t lo = 0, hi = T->n_scan;
    while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (T->scan_key[mid] < key) lo = mid + 1; else hi = mid; }
    return (lo < T->n_scan && T->scan_key[lo] == key) ? lo : -1;
}

int64_t mtdo_contact_tracing(const mtdo_params *P, const mtdo_dayctx *D, const mtdo_trace_tape *T, int64_t n_traced,
                             const int32_t *traced, int64_t n_slices, const int64_t *slice_off, const int32_t *rows3,
                             int64_t n_res_cells, const int64_t *res_off, const int32_t *res_nres,
                             const int32_t *res_members, mtdo_agents *A, const uint8_t *epi_present, int32_t n_locals,
                             uint8_t *primary, uint8_t *secondary, mtdo_trace_stats *S)
{
    const int64_t N = A->n_agents;
    uint8_t *is_traced = (uint8_t *)calloc((size_t)N + 1, 1);
    for (int64_t i = 0; i < n_traced; i++) if (traced[i] >= 0 && traced[i] < N) is_traced[traced[i]] = 1;
    memset(primary, 0, (size_t)N); memset(secondary, 0, (size_t)N);
    memset(S, 0, sizeof(*S));

    /* ---- scan of the look-back slices (:677-721) ---- */
    for (int64_t back = 0; back < n_slices; back++) {
        for (int64_t r = slice_off[back]; r < slice_off[back + 1]; r++) {
            int sct = rows3[3 * r], a1 = rows3[3 * r + 1], a2 = rows3[3 * r + 2];
            int f1 = is_traced[a1], f2 = is_traced[a2];
            if (f1 == f2) continue;                                         /* exactly one endpoint is a traced positive */
            double u;
            uint64_t idx = (uint64_t)(r - slice_off[back]);
            if (D->mode == MTDO_MODE_REPLAY) {
                int64_t t = trace_find_scan(T, ((uint64_t)back << 32) | idx);
                if (t < 0) { S->tape_miss++; continue; }
                u = T->scan_u[t];
            } else {
                uint32_t w[4];
                philox_site(D, MTDO_SITE_TRACE, (uint32_t)back, (uint32_t)a1, (uint32_t)a2, w);   /* keyed by the row's pair, not by its
                                                                                       * index: the same draw however the rows are sharded */
                u = mtdo_u53(w[0], w[1]);
            }
            S->n_scan_draws++;
            if (u < P->tracing_success_prob[sct]) primary[f1 ? a2 : a1] = 1;    /* :714-721 */
        }
    }

    /* ---- processing order of the primary contacts: the reference iterates a Python set (:723); replay takes the
     * recorded order, Philox mode uses ascending agent id ---- */
    int64_t np_ = 0;
    for (int64_t a = 0; a < N; a++) np_ += primary[a];
    int32_t *order = (int32_t *)malloc(sizeof(int32_t) * (size_t)(np_ + 1));
    if (D->mode == MTDO_MODE_REPLAY) {
        int64_t k = 0;
        for (int64_t i = 0; i < T->n_prim && k < np_; i++) {
            if (primary[T->prim_id[i]]) order[k++] = T->prim_id[i]; else S->tape_miss++;
        }
        /* primaries without an epi record are skipped by the reference before any draw (:729-733): not on the tape */
        for (int64_t a = 0; a < N && k < np_; a++) {
            if (!primary[a]) continue;
            int seen = 0;
            for (int64_t i = 0; i < T->n_prim; i++) if (T->prim_id[i] == a) { seen = 1; break; }
            if (!seen) { if (epi_present && !epi_present[a]) order[k++] = (int32_t)a; else { S->tape_miss++; order[k++] = (int32_t)a; } }
        }
    } else {
        int64_t k = 0;
        for (int64_t a = 0; a < N; a++) if (primary[a]) order[k++] = (int32_t)a;
    }

    uint8_t *res_done = (uint8_t *)calloc((size_t)n_res_cells + 1, 1);
    int32_t *sel = NULL; int64_t sel_cap = 0;
    for (int64_t oi = 0; oi < np_; oi++) {
        const int32_t p = order[oi];
        if (epi_present && !epi_present[p]) continue;                      /* agents_epi[contact_id] missing / empty */
        tracedraws td = { D, P, p, NULL, 0, 0, S };
        if (D->mode == MTDO_MODE_REPLAY) {
            int64_t i = 0;
            while (i < T->n_prim && T->prim_id[i] != p) i++;
            if (i < T->n_prim) { td.rec = T->prim_draws + T->prim_off[i]; td.len = T->prim_off[i + 1] - T->prim_off[i]; }
        }
        int ts = td_ts(&td, 0);
        trace_quarantine(&td, A, p, D->day, ts, Q_POSITIVE_CONTACT, 1);
        ts = td_ts(&td, 2);
        trace_test(&td, A, p, D->day, ts, Q_POSITIVE_CONTACT, 3);
        /* secondary contacts: once per residence, triggered by its first primary contact (:737-817) */
        int32_t res = A->res_cellid[p];
        if (res < 0 || res >= n_res_cells || res_done[res]) { if (D->mode == MTDO_MODE_REPLAY && td.pos != td.len) S->tape_miss++; continue; }
        res_done[res] = 1;
        int64_t L = 0, b = res_off[res], e = res_off[res + 1];
        if (e - b > sel_cap) { sel_cap = (e - b) * 2 + 16; sel = (int32_t *)realloc(sel, sizeof(int32_t) * (size_t)sel_cap); }
        for (int64_t q = b; q < e; q++) {
            int32_t m = res_members[q];
            if (q - b < res_nres[res] && m == p) continue;                 /* np.delete(resident_ids, == contact_id) */
            sel[L++] = m;
        }
        if (L == 0) { if (D->mode == MTDO_MODE_REPLAY && td.pos != td.len) S->tape_miss++; continue; }
        int64_t k = (int64_t)py_round((double)L * P->tracing_success_prob[0]);
        if (k > L) k = L;
        for (int64_t t = 0; t < k; t++) {                                  /* np.random.choice(sec, k, replace=False) */
            if (D->mode == MTDO_MODE_REPLAY) {
                int32_t id = (int32_t)td_next(&td);
                int64_t r = t; while (r < L && sel[r] != id) r++;
                if (r == L) { S->tape_miss++; r = t; }
                int32_t y = sel[r]; for (int64_t j = r; j > t; j--) sel[j] = sel[j - 1]; sel[t] = y;
            } else {
                uint32_t w[4];
                philox_site(D, MTDO_SITE_TRACE_PRIM, (uint32_t)p, (uint32_t)(16 + t), 0, w);
                int64_t r = t + (int64_t)(((uint64_t)w[0] * (uint64_t)(L - t)) >> 32);
                int32_t y = sel[r]; for (int64_t j = r; j > t; j--) sel[j] = sel[j - 1]; sel[t] = y;
            }
        }
        for (int64_t j = 0; j < k; j++) {
            int32_t s = sel[j];
            if (s < 0 || s >= N || secondary[s] || primary[s]) continue;   /* skip if already handled (:798) */
            /* quirk B9: the PRIMARY's record is the one offered to schedule_quarantine / schedule_test (:800-809) */
            uint32_t base = 65536u + 8u * (uint32_t)j;
            int t1 = td_ts(&td, base);
            trace_quarantine(&td, A, p, D->day, t1, Q_SECONDARY_CONTACT, base + 1);
            int t2 = td_ts(&td, base + 2);
            trace_test(&td, A, p, D->day, t2, Q_SECONDARY_CONTACT, base + 3);
            secondary[s] = 1;
        }
        if (D->mode == MTDO_MODE_REPLAY && td.pos != td.len) S->tape_miss++;
    }
    for (int64_t a = 0; a < N; a++) { S->n_primary += primary[a]; S->n_secondary += secondary[a]; }
    free(is_traced); free(order); free(res_done); free(sel);
    return S->n_primary + S->n_secondary;
}
