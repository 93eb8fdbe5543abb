/*
 * include/mtd_b200.h -- C ABI of libmtd_b200.so: the B200 (sm_100a) implementation of mtdcovabm's per-day
 * contact-network + transmission + SEIR-advance path.  Plain C: pointers and sizes only, no torch / C++ types.
 *
 * Every entry point replaces a reference interface (paths relative to /root/reference/simulator):
 *   mtd_simulate_contact_network   ContactNetwork.simulate_contact_network        contactnetwork.py:65-124
 *                                  (+ generate_contact_network                    contactnetwork.py:134-315)
 *   mtd_simulate_direct_contacts   Epidemiology.simulate_direct_contacts          epidemiology.py:88-231
 *                                  (+ simulate_seir_state_transition, schedule_*  epidemiology.py:233-660)
 *   mtd_advance_seir               seirstateutil.update_agent_state               seirstateutil.py:38-72
 *   mtd_set_static                 Static.get(id, "age_bracket_index" | "epi_age_bracket_index" | "soc_rate" |
 *                                  "res_cellid")                                  static.py:351-363
 *   mtd_set_state / mtd_get_state  vars_util.agents_seir_state / agents_infection_type / agents_infection_severity /
 *                                  agents_vaccination_doses (vars.py:52-56) and agents_epi[id][...] (agentsutil.py:56-62)
 *   mtd_set_params                 Epidemiology.__init__ / ContactNetwork.__init__ parameter extraction
 *                                  epidemiology.py:29-70, contactnetwork.py:41-42
 *   mtd_get_contacts               vars_util.directcontacts_by_simcelltype_by_day rows   contactnetwork.py:98-105
 *   mtd_apply_deltas               the "updated agents" exchange of the reference's stateful-actor strategy
 *                                  actor_dist.py:491-518 (multi-GPU: one allgather per day in between)
 *
 * Conventions: every function returns 0 on success, non-zero MTD_E_* otherwise; mtd_last_error(h) gives the text.
 * No C++ exception crosses this boundary.  One handle = one device + one stream; calls on a handle are not
 * re-entrant.  `mem` arguments say where the caller's buffers live: MTD_MEM_HOST (the library copies through its
 * stream) or MTD_MEM_DEVICE (used in place / copied device-to-device).
 */
#ifndef MTD_B200_H
#define MTD_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MTD_MAXSCHED 12          /* live + stale schedule entries per agent */
#define MTD_NONE (-32768)        /* "None" for day-valued record fields */
#define MTD_EXPO_W 21            /* doubles per replay-tape exposure record (tests/golden/ref_harness.py) */
#define MTD_DELTA_WORDS 20       /* int32 words per exposure delta record */
#define MTD_MAX_CELL_INTERVALS 16384 /* largest cell (intervals of one cell on one day); cells up to 1024 intervals are processed
                                      * in shared memory, larger ones in a global workspace sized from max_intervals */

enum { MTD_OK = 0, MTD_E_ARG = 1, MTD_E_CUDA = 2, MTD_E_CAPACITY = 3, MTD_E_CELL_TOO_LARGE = 4, MTD_E_STATE = 5, MTD_E_TIMEOUT = 6 };
enum { MTD_MEM_HOST = 0, MTD_MEM_DEVICE = 1, MTD_MEM_STAGED = 2 /* the stream given to mtd_prefetch_intervals */,
       MTD_MEM_BUILT = 3 /* the stream built on the device by mtd_build_intervals */ };
enum { MTD_MODE_REPLAY = 0, MTD_MODE_PHILOX = 1 };

typedef struct mtd_handle_s *mtd_handle;

typedef struct {
    int32_t device;              /* CUDA device ordinal */
    int64_t n_agents;            /* N: agent ids are dense in [0, N) */
    int64_t n_cells;             /* cell ids are dense in [0, n_cells) */
    int64_t max_intervals;       /* capacity M of one day's interval stream */
    int64_t max_rows;            /* capacity of the dense contact-row buffer of one day (0 = derive from M) */
    int64_t max_exposures;       /* capacity of the per-day exposure delta buffer (0 = derive) */
    void *stream;                /* cudaStream_t to run on; NULL = library-owned stream */
} mtd_config;

/* Same field order as oracle/mtd_oracle.h:mtdo_params (kept separate on purpose: the oracle is not product code). */
typedef struct {
    double cm[16][6];
    int32_t k_by_act[6];
    double p_household, p_workplace, p_school, p_community;
    double probs_by_age[5][11];
    double ln_mu[10], ln_sigma[10];
    double test_mu, test_sigma, result_mu, result_sigma;
    double testing_after_symptoms_probability;
    double quarantine_after_symptoms_probability;
    double ignore_quarantine_rules_probability;
    int32_t quarantine_positive_duration, quarantine_positive_contact_duration, quarantine_secondary_contact_duration;
    int32_t contact_tracing_days_back, contact_tracing_positive_delay_days, contact_tracing_secondary_delay_days;
    double tracing_success_prob[4];
    double vacc_immunity, vacc_asymptomatic, vacc_decay_interval;
    double testing_false_positive_rate, testing_false_negative_rate;   /* epidemiology.py:59-60 */
} mtd_params;

typedef struct {
    int32_t mode;                /* MTD_MODE_* */
    uint32_t seed;               /* Philox key word 0 */
    int32_t day;
    double mhd;                  /* dyn_params.masks_hygiene_distancing_multiplier (dynamicparams.py:67-73) */
    int32_t testing_enabled, quarantine_enabled;
} mtd_day;

/* replay tape of one day (host pointers; sorted by key) */
typedef struct {
    /* selections: key = cell << 32 | main agent; sel_picks[sel_off[i] .. sel_off[i+1]) = the partner AGENT IDS the
     * reference selected for that main agent, in order (contactnetwork.py:299) */
    int64_t n_sel;  const uint64_t *sel_key;  const int32_t *sel_off;  const int32_t *sel_picks;
    int64_t n_pair; const uint64_t *pair_k1;  const uint32_t *pair_k2; const double *pair_u;
    int64_t n_expo; const int32_t *expo_agent; const double *expo_row;
} mtd_tape;

typedef struct {
    int64_t n_slots;             /* active cells of the day */
    int64_t n_rows;              /* direct contacts (dense rows) */
    int64_t n_exposed;           /* exposures applied */
    int64_t pairs_evaluated;     /* sum over cells of #{i<j: different agents}: the headline metric's unit */
    int64_t n_potential;         /* overlapping agent pairs */
    int64_t n_si_rows;           /* rows with exactly one Susceptible and one Infectious endpoint */
    int64_t sched_overflow;      /* schedule entries dropped (MTD_MAXSCHED reached) */
    int64_t tape_miss;           /* replay lookups without a record */
    int64_t class_cells[9];      /* cells per size class (<=8, <=16, <=32, <=64, <=128, <=192, <=320, <=1024 intervals, above) */
    int64_t n_incidences;        /* sum over cells of distinct agents (cell, agent incidences) */
    int64_t error;               /* MTD_E_* raised on the device (capacity, cell too large) */
} mtd_counters;

int mtd_create(const mtd_config *cfg, mtd_handle *out);
int mtd_destroy(mtd_handle h);
const char *mtd_last_error(mtd_handle h);
int mtd_abi_version(void);

int mtd_set_params(mtd_handle h, const mtd_params *p);
int mtd_set_cells(mtd_handle h, int64_t n_cells, const int8_t *cell_type, const int16_t *cell_indid, int mem);
int mtd_set_static(mtd_handle h, int64_t n, const int8_t *age_bracket, const int8_t *epi_age_bracket,
                   const double *soc_rate, const int32_t *res_cellid, int mem);
/* any pointer may be NULL = leave that array untouched */
int mtd_set_state(mtd_handle h, int64_t n, const int8_t *seir, const int8_t *inf_type, const int8_t *inf_sev,
                  const int32_t *last_dose_day, const int32_t *sched, const uint8_t *sched_cnt,
                  const int32_t *test_day, const int32_t *test_result_day, const int32_t *quarantine,
                  const int32_t *hosp, int mem);
int mtd_get_state(mtd_handle h, int64_t n, int8_t *seir, int8_t *inf_type, int8_t *inf_sev, int32_t *sched,
                  uint8_t *sched_cnt, int32_t *test_day, int32_t *test_result_day, int32_t *quarantine,
                  int32_t *hosp, int mem);
int mtd_set_tape(mtd_handle h, const mtd_tape *tape);

/* K8: interval builder = Itinerary.update_cell_agents_timesteps + combine_same_cell_itinerary_entries
 * (itinerary.py:1403-1488), the producer of this path's input.  One record per itinerary: its entries
 * {start timestep -> cell} sorted by timestep (CSR ent_off[n_itin + 1] / ent_ts / ent_cell; cell -1 = residence) and
 * the agents sharing it (CSR grp_off[n_itin + 1] / grp_agent / grp_rescell; a local agent is a group of one).  Builds
 * the day's interval stream on the device, in the reference's emission order; n_intervals receives its length.  Run the
 * day on it with mtd_simulate_contact_network(h, day, n_intervals, NULL, NULL, NULL, NULL, NULL, MTD_MEM_BUILT);
 * mtd_get_intervals copies the stream out. */
int mtd_build_intervals(mtd_handle h, int64_t n_itin, const int64_t *ent_off, const int16_t *ent_ts, const int32_t *ent_cell,
                        const int64_t *grp_off, const int32_t *grp_agent, const int32_t *grp_rescell, int mem,
                        int64_t *n_intervals);
int mtd_get_intervals(mtd_handle h, int64_t n, int32_t *iv_agent, int32_t *iv_cell, uint8_t *iv_start, uint8_t *iv_end, int mem);

/* K0 (CSR build) + K1 (pairs, degrees, greedy sampling) + K2 (dense rows).  Interval stream in emission order
 * (itinerary.py:1446-1452); iv_order may be NULL (= position in the stream) or strictly increasing global
 * emission indices when the stream is one rank's shard. */
int mtd_simulate_contact_network(mtd_handle h, const mtd_day *day, int64_t n_intervals, const int32_t *iv_agent,
                                 const int32_t *iv_cell, const uint8_t *iv_start, const uint8_t *iv_end,
                                 const uint32_t *iv_order, int mem);
/* K3: pair evaluation, winner resolution and trajectory sampling into fixed-size delta records, then (apply != 0)
 * application to the per-agent state.  delta_buf is a DEVICE buffer of (1 + cap_records) records of
 * MTD_DELTA_WORDS int32 each -- record 0 is a header whose word 0 receives the record count -- or NULL to use the
 * handle's own buffer.  Multi-GPU: pass this rank's slice of the allgather buffer with apply == 0, allgather, then
 * mtd_apply_deltas over all n_bufs slices. */
int mtd_simulate_direct_contacts(mtd_handle h, const mtd_day *day, int32_t *delta_buf, int64_t cap_records, int apply);
/* K9: Epidemiology.simulate_seir_state_transition(current_day, agent, id, incremental_days, overlapping_timesteps, ...)
 * called on its own (epidemiology.py:233-500) -- the itinerary stage's initial cases: on simulated day 1 every agent that
 * starts Infectious is given an exposure day in [day - pre_exposure_days, day) and its whole trajectory is sampled
 * from there (itinerary.py:373-395), with the states reset to those preceding the first transition still ahead
 * (reset_states, epidemiology.py:243).  n DISTINCT agent ids; exposure_day[n] or NULL (drawn on the device, Philox mode
 * only); timesteps[n_timesteps] = the overlapping_timesteps list shared by the batch (the itinerary passes
 * arange(144)); asym_mult[n] or NULL (1.0, the reference's default); draws: replay mode, n x MTD_EXPO_W doubles in the
 * slot layout of the tape's exposure records, else NULL; out8 (optional): 8 int32 per case = [seir state, infection
 * type, severity, recovered, schedule entries appended, exposure timestep, exposure day, symptomatic day or -1].
 * HOST pointers; synchronous.  Schedule entries, states, test and hospitalisation records go to the handle's state. */
int mtd_seed_exposures(mtd_handle h, const mtd_day *day, int64_t n, const int32_t *agent_ids, const int32_t *exposure_day,
                       int32_t pre_exposure_days, const uint8_t *timesteps, int32_t n_timesteps, const double *asym_mult,
                       const double *draws, int32_t *out8);
/* Epidemiology.simulate_direct_contacts(agents_directcontacts, cellid, cell_type, day) called on its own
 * (epidemiology.py:88): the caller's contact pairs of ONE cell with their overlap lists [(start, end), ...]
 * (ov_off[n_rows + 1] offsets into ov_se, two bytes per overlap) become the day's dense rows, in the order given;
 * mtd_simulate_direct_contacts then evaluates them exactly like rows produced by mtd_simulate_contact_network. */
int mtd_set_contacts(mtd_handle h, const mtd_day *day, int32_t cellid, int64_t n_rows, const int32_t *a1,
                     const int32_t *a2, const int64_t *ov_off, const uint8_t *ov_se, int mem);
int mtd_apply_deltas(mtd_handle h, const int32_t *delta_bufs, int64_t n_bufs, int64_t cap_records);

/* Multi-GPU exchange of the exposure records over NVLink peer memory (replaces the reference's
 * actor-to-actor send_results / receive_results of updated agents, actor_dist.py:491-518, and the per-day NCCL
 * all-gather of fixed-capacity buffers).  One process per GPU on one box:
 *   mtd_exchange_init   allocate this rank's exchange buffer (two day-parity sets of `world` slices of cap_records
 *                       records) and return its CUDA IPC handle (64 bytes) for the peers;
 *   mtd_exchange_open   map a peer's buffer from its IPC handle (every peer, once);
 *   mtd_exchange_ready  check that every peer is mapped and enable the path (callers barrier after this).
 * mtd_simulate_direct_contacts(h, day, NULL, 0, MTD_APPLY_EXCHANGE) then writes the day's records into the rank's own
 * slice, pushes exactly those records into the same slice of every peer with stores over NVLink, raises an arrival
 * flag per peer, waits (on the device, with a 10 s timeout -> MTD_E_TIMEOUT) for every peer's flag and applies all
 * slices in rank order with the same earliest-key rule as the single-GPU path.  No host synchronisation, no NCCL. */
#define MTD_APPLY_EXCHANGE 2
#define MTD_IPC_HANDLE_BYTES 64
int mtd_exchange_init(mtd_handle h, int rank, int world, int64_t cap_records, void *ipc_handle_out);
int mtd_exchange_open(mtd_handle h, int peer, const void *ipc_handle);
int mtd_exchange_ready(mtd_handle h);
/* K4: daily SEIR advance for every agent with a schedule entry dated `day`; mask (uint8[N]) may be NULL. */
int mtd_advance_seir(mtd_handle h, int32_t day, const uint8_t *mask, int mem);
/* Intervention coupling for day loops that do not run the reference's itinerary stage (tiled configurations, SURVEY.md
 * section 8f rank 1; the stage's quarantine / hospital / deceased handling is itinerary.py:2283-2528): when on, K0
 * drops every interval of an agent who today is deceased, inside its quarantine window or inside its hospitalisation
 * window, unless the interval is in the agent's residence cell. */
int mtd_set_intervention_mask(mtd_handle h, int on);
/* Day pipelining (no reference counterpart: the reference runs its stages one after the other, sim.py:1196-1365).
 * K0 of a day depends on the day's interval stream and on static cell attributes only, K1 on K0's products and static
 * agent attributes, so the library runs K0 and K1 of the coming days on streams of their own beside K2/K3 of the
 * current day (K0's products, K1's scratch rows and the day's counters are double-buffered by day parity).  For
 * MTD_MEM_DEVICE inputs the library cannot see who wrote the arrays: with `on`, the caller promises that the interval
 * arrays passed to mtd_simulate_contact_network are complete when the call is made (no pending work of any stream
 * still writes them) and stay untouched until the call after the next one.  Off by default; results are identical
 * either way.  Not used while the intervention mask is on (K0 then reads the day's state). */
int mtd_set_k0_overlap(mtd_handle h, int on);

/* K7: vaccination.
 *   mtd_enable_vaccination     allocate the per-agent [day, timestep] vaccination record (agent["vaccination_days"][-1],
 *                              epidemiology.py:961-967); locals are the ids [0, n_locals).  From then on
 *                              mtd_advance_seir also does the dose bookkeeping of the vaccination day
 *                              (itinerary.py:2489-2527): record dated today -> last_dose_day = today
 *                              (agents_vaccination_doses[a].append(day)); record dated in the past -> cleared.
 *   mtd_schedule_vaccinations  Epidemiology.schedule_vaccinations (epidemiology.py:911-976), device part: k =
 *                              dyn_params.num_vaccinations_today locals without a dose are drawn without replacement
 *                              (the k smallest of unique Philox keys, radix select) and get [day + 1, timestep in
 *                              42..120].  Replay mode: the reference's own sample (ids + timesteps).  selected: optional
 *                              u8[N] flags of the agents scheduled by this call; n_selected: their number.
 *   mtd_set_vaccination / mtd_get_vaccination   last_dose_day i32[N] and the records i32[N][2] (MTD_NONE = none). */
int mtd_enable_vaccination(mtd_handle h, int64_t n_locals);
int mtd_schedule_vaccinations(mtd_handle h, const mtd_day *day, int64_t k, int64_t n_replay, const int32_t *replay_ids,
                              const int32_t *replay_ts, uint8_t *selected, int64_t *n_selected, int mem);
int mtd_set_vaccination(mtd_handle h, int64_t n, const int32_t *last_dose_day, const int32_t *vacc_day, int mem);
int mtd_get_vaccination(mtd_handle h, int64_t n, int32_t *last_dose_day, int32_t *vacc_day, int mem);

/* results of the last day step */
int mtd_get_counters(mtd_handle h, mtd_counters *out);           /* synchronises the stream */
int mtd_get_contacts(mtd_handle h, int32_t *rows5, int32_t *row_dur, int32_t *row_cell, int64_t cap, int64_t *n,
                     int mem);                                    /* rows5: [n][5] simcelltype,a1,a2,start,end */
int mtd_get_exposed(mtd_handle h, int32_t *agent_ids, int64_t cap, int64_t *n, int mem);   /* in reference order */
int mtd_get_infection_prob(mtd_handle h, double *p, int64_t cap, int mem);   /* per dense row, NaN if not S-I */
int mtd_occupancy(mtd_handle h, int64_t out[6]);                 /* dynamicstatistics.py:169-193 */
int mtd_sync(mtd_handle h);

/* ---- K5: contact tracing (Epidemiology.contact_tracing, epidemiology.py:667-834) ----
 * replay tape of one call (host pointers): scan uniforms keyed by (look-back distance << 32 | row index in its slice),
 * sorted; per primary contact, in the reference's processing order, the ordered draws it consumed. */
typedef struct {
    int64_t n_scan; const uint64_t *scan_key; const double *scan_u;
    int64_t n_prim; const int32_t *prim_id; const int32_t *prim_off; const double *prim_draws;
} mtd_trace_tape;
typedef struct { int64_t n_primary, n_secondary, n_scan_draws, tape_miss; } mtd_trace_stats;
/* residences (cells_households / cells_institutions / cells_accommodation, epidemiology.py:752-788): CSR over cell ids,
 * members = residents (the first res_nres[c]) followed by staff */
int mtd_set_residences(mtd_handle h, int64_t n_res_cells, const int64_t *res_off, const int32_t *res_nres,
                       const int32_t *res_members, int mem);
/* traced: vars_util.contact_tracing_agent_ids.  rows3 [n][3] = simcelltype, a1, a2 of the look-back slices the caller
 * cut from directcontacts_by_simcelltype_by_day (slice s = s days back; slice_off is a HOST array of n_slices + 1).
 * epi_present (uint8[N], may be NULL = everybody): agents_epi[id] exists.  Updates the quarantine / test records of
 * the primary contacts in the handle's state. */
int mtd_contact_tracing(mtd_handle h, const mtd_day *day, const mtd_trace_tape *tape, int64_t n_traced,
                        const int32_t *traced, int64_t n_slices, const int64_t *slice_off, const int32_t *rows3,
                        const uint8_t *epi_present, int mem);
int mtd_get_traced(mtd_handle h, uint8_t *primary, uint8_t *secondary, mtd_trace_stats *stats, int mem);
/* Cell-sharded contact tracing (the reference's stateful strategy syncs the traced agents between the partitions at the
 * end of the day, sim.py:2034-2063, actor_dist.py:432-569).  The contact rows stay on the rank that produced them:
 *   mtd_contact_tracing_scan    marks the traced positives and scans THIS rank's look-back slices; primary_out (device
 *                               u8[N]) receives the primary-contact flags found in them.  Philox mode only: the scan draw
 *                               is keyed by (day, look-back distance, a1, a2), so it does not depend on the sharding;
 *   (caller)                    OR / MAX all-reduce of the flags over the ranks -- the path's second exchange step;
 *   mtd_contact_tracing_finish  takes the merged flags (device u8[N]) and runs the scheduling phases on this rank's
 *                               replica of the per-agent records: identical on every rank, identical to one GPU. */
int mtd_contact_tracing_scan(mtd_handle h, const mtd_day *day, int64_t n_traced, const int32_t *traced, int64_t n_slices,
                             const int64_t *slice_off, const int32_t *rows3, const uint8_t *epi_present,
                             uint8_t *primary_out, int mem);
int mtd_contact_tracing_finish(mtd_handle h, const uint8_t *primary_in);
/* K6: the record bookkeeping of Itinerary.sample_intervention_activities (itinerary.py:2283-2528; SURVEY.md section 8f
 * rank 4), once per day over all agents (epi_present: optional u8[N], agents without an epi record are skipped), in the
 * reference's order: (A) :2301-2332 results due today flag the agent positive (false-negative / false-positive rates of
 * epidemiology.py:59-60), positive agents get a Positive quarantine (epidemiology.py:568-625) and are returned,
 * ascending, as the day's contact_tracing_agent_ids -- the `traced` input of mtd_contact_tracing; result days in the past
 * are cleared; (B) :2396 hospitalisation_days cleared on their end day; (C) :2456 quarantine_days cleared on their end
 * day (not when the window starts today, nor when the hospitalisation ended today inside it); (D) :2486-2487 test_day
 * cleared six days after the test.  Call it AFTER mtd_advance_seir of the day: the reference advances the agent's state
 * first (itinerary.py:399,1587) and judges the result on the new state.  The stage's re-timing of records on arrival /
 * departure days belongs to the itinerary and is not done here.
 * Draws: Philox keyed (day, agent); replay mode: the tape of mtd_set_k6_tape (HOST pointers: strictly ascending agent
 * ids, two doubles per agent = the result draw of :2310/:2322 and the quarantine draw of epidemiology.py:580). */
int mtd_set_k6_tape(mtd_handle h, int64_t n, const int32_t *agent_ids, const double *draws2);
int mtd_process_test_results(mtd_handle h, const mtd_day *day, const uint8_t *epi_present, int32_t *traced_ids,
                             int64_t cap, int64_t *n, int mem);

/* ---- pipelined boundary (the day loop of sim.py:1392-2125 driven from host memory without stalling the device) ----
 * The reference hands the stage host containers and gets host containers back (contactnetwork.py:65-124).  These three
 * calls keep that contract -- pinned HOST buffers in, pinned HOST buffers out -- while the copies run beside the
 * kernels: day d+1's intervals are uploaded by a copy stream during day d, and day d's results are copied into the
 * caller's pinned memory by the copy engine on a side stream during day d+1 (no SM is taken from the kernels).
 *   mtd_prefetch_intervals    enqueue the upload of the NEXT day's stream (pinned host pointers); the next
 *                             mtd_simulate_contact_network call passes mem = MTD_MEM_STAGED (pointers ignored).
 *   mtd_export_day_async      enqueue the export of the day just computed; needs mtd_set_export(h, 1) before that
 *                             day's mtd_simulate_contact_network.  rows12: packed contact rows, 12 bytes each
 *                             {int32 a1, int32 a2, uint8 start, uint8 end, uint8 simcelltype, uint8 0} in the
 *                             reference's row order (contactnetwork.py:98-105); exposed_kv: n_exposed pairs
 *                             {uint64 order key, uint64 agent id} (sort by key for the reference's order);
 *                             result: counts + occupancy.  mem = MTD_MEM_HOST: the buffers are pinned (page-locked)
 *                             host memory; the result record is copied first, and the two bulk copies are issued
 *                             with their exact sizes as soon as it has landed -- by mtd_export_wait, or by the next
 *                             call that reuses the export slot (two days later).  MTD_MEM_DEVICE: device memory.
 *   mtd_export_wait           block the host until the export with that ticket has landed (all three buffers). */
typedef struct {
    int64_t day, n_rows, n_exposed;
    int64_t occupancy[6];            /* agents per SEIR state after the day (dynamicstatistics.py:169-193) */
    int64_t pairs_evaluated, error, ticket;
} mtd_day_result;
int mtd_set_export(mtd_handle h, int on);
int mtd_prefetch_intervals(mtd_handle h, int64_t n_intervals, const int32_t *iv_agent, const int32_t *iv_cell,
                           const uint8_t *iv_start, const uint8_t *iv_end, const uint32_t *iv_order);
int mtd_export_day_async(mtd_handle h, void *rows12, int64_t cap_rows, uint64_t *exposed_kv, int64_t cap_exposed,
                         mtd_day_result *result, int mem, int64_t *ticket);
int mtd_export_wait(mtd_handle h, int64_t ticket);

/* Measurement hooks (the reference's per-stage wall-clock columns, sim.py:231-244,1852-1856): with profiling on,
 * CUDA events are recorded on the handle's stream around every stage of the day step; mtd_get_stage_ms
 * synchronises and returns, per stage, the MEAN device time over its executions since the previous call (the first
 * 512 of them) and the number of kernels the library launched since the previous call. */
enum { MTD_STAGE_H2D = 0, MTD_STAGE_K0 = 1, MTD_STAGE_K1 = 2, MTD_STAGE_K2 = 3, MTD_STAGE_K3_EVAL = 4,
       MTD_STAGE_K3_APPLY = 5, MTD_STAGE_K4 = 6, MTD_N_STAGES = 7 };
int mtd_set_profiling(mtd_handle h, int on);
/* test hook: out[i] = CUDA's fp64 log(x[i]) (HOST pointers), the function K1 uses for its two per-cell logarithms
 * log(avg_ts), log(avg_pc) (contactnetwork.py:240-245 uses glibc's through math.log); tests/test_device_log.py */
int mtd_debug_log(mtd_handle h, int64_t n, const double *x, double *out);
int mtd_get_stage_ms(mtd_handle h, float ms[MTD_N_STAGES], int64_t launches[MTD_N_STAGES]);

/* Occupancy count without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same result as mtd_occupancy (agents per SEIR state 0..5), but handle-free: the caller passes the input
 * (seir: DEVICE int8[n_agents]) and the output (out6: DEVICE int64[6]) and the stream to run on (NULL = the default
 * stream); the call is asynchronous on that stream.  The kernel's scratch (per-block partial counts + an arrival
 * ticket) is a file-scope __device__ array of MTD_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until
 * exit; the entry point calls no cudaMalloc / cudaFree / allocating library routine.  The array's device address is
 * resolved once per device on the first call for that device (make that call outside a graph capture).
 * n_agents > MTD_NOALLOC_MAX_AGENTS -> MTD_E_CAPACITY with the text in err (err / err_cap: optional caller buffer for
 * the error message).  All calls on a device share the array: calls that overlap on one GPU race -- serialise them. */
#define MTD_NOALLOC_MAX_AGENTS (1LL << 26)         /* 67 108 864 agent ids: 1024 blocks of 65536 */
#define MTD_NOALLOC_SCRATCH_BYTES 65792            /* 1024 rows x 8 x u64 + 256 bytes (ticket) */
int mtd_occupancy_noalloc(int32_t device, const int8_t *seir, int64_t n_agents, int64_t *out6, void *stream,
                          char *err, int64_t err_cap);

#ifdef __cplusplus
}
#endif
#endif
