// K7 -- vaccination scheduling (Epidemiology.schedule_vaccinations, epidemiology.py:911-976) and the dose bookkeeping
// of the vaccination day (itinerary.py:2489-2526).
//
// schedule: the reference draws `k = num_vaccinations_today` agents without replacement from the locals that have no
// dose yet (np.random.choice over set(all) ^ set(vaccinated), :951-958) and gives each a [day + 1, timestep] record with
// a uniform timestep from arange(42, 121) (:963-965).  Here every candidate gets a unique 64-bit key
// (Philox word << 32 | agent id); the k smallest keys are the sample -- a uniformly random k-subset, the same
// distribution -- found by a most-significant-digit radix SELECT: four 16-bit histogram passes narrow the key prefix to
// the k-th smallest key, one marking pass applies it.  No sort, no host round trip, exact k.
// dose day: on its vaccination day an agent's dose list grows by today (last_dose_day = day, :2517-2522; the
// travel-rescheduling branch belongs to the itinerary stage and is not part of this path); a record dated in the past
// is cleared (:2526-2527).
#include "common.cuh"

namespace {

struct SelState {                      // device-side state of the radix select
    unsigned long long prefix;         // key prefix fixed so far (high bits)
    long long remaining;               // rank still to resolve inside the prefix bucket
    long long n_candidates;
    long long k_eff;
};

__device__ __forceinline__ bool vacc_candidate(const DevCtx &c, int64_t a) { return c.last_dose_day[a] == MTD_NONE; }

__device__ __forceinline__ unsigned long long vacc_key(const DevCtx &c, int64_t a, uint32_t *w1)
{
    uint32_t w[4];
    philox4x32((uint32_t)c.day.day, (uint32_t)a, 0u, 0u, c.day.seed, 7u, w);
    if (w1) *w1 = w[1];
    return ((unsigned long long)w[0] << 32) | (unsigned long long)(uint32_t)a;
}

__global__ void k7_count_kernel(DevCtx c, int64_t n_locals, SelState *S)
{
    long long n = 0;
    for (int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; a < n_locals; a += (int64_t)gridDim.x * blockDim.x)
        n += vacc_candidate(c, a);
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(FULLMASK, n, o);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd((unsigned long long *)&S->n_candidates, (unsigned long long)n);
}

// histogram of digit `pass` (0 = most significant 16 bits) over the candidates whose higher digits equal the prefix
__global__ void k7_hist_kernel(DevCtx c, int64_t n_locals, const SelState *S, int pass, unsigned int *hist)
{
    const int shift = 48 - 16 * pass;
    const unsigned long long prefix = S->prefix;
    if (S->k_eff <= 0 || S->k_eff >= S->n_candidates) return;      // nothing / everybody: no selection needed
    for (int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; a < n_locals; a += (int64_t)gridDim.x * blockDim.x) {
        if (!vacc_candidate(c, a)) continue;
        const unsigned long long key = vacc_key(c, a, nullptr);
        if (pass > 0 && (key >> (shift + 16)) != (prefix >> (shift + 16))) continue;
        atomicAdd(&hist[(key >> shift) & 0xffffu], 1u);
    }
}

// one block: the bucket holding the remaining-th smallest key -> extend the prefix
__global__ void k7_init_kernel(SelState *S, long long k)
{
    const long long ke = k < 0 ? 0 : (k > S->n_candidates ? S->n_candidates : k);
    S->k_eff = ke; S->remaining = ke - 1; S->prefix = 0;
}

__global__ void k7_pick_kernel(SelState *S, int pass, unsigned int *hist)
{
    __shared__ long long part[256];
    if (S->k_eff <= 0 || S->k_eff >= S->n_candidates) return;
    const int shift = 48 - 16 * pass;
    long long s = 0;
    for (int b = threadIdx.x * 256; b < threadIdx.x * 256 + 256; b++) s += hist[b];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        long long rem = S->remaining, acc = 0;
        int t = 0;
        while (t < 255 && acc + part[t] <= rem) { acc += part[t]; t++; }
        int b = t * 256;
        while (b < t * 256 + 255 && acc + hist[b] <= rem) { acc += hist[b]; b++; }
        S->remaining = rem - acc;
        S->prefix |= (unsigned long long)b << shift;
    }
    __syncthreads();
    for (int b = threadIdx.x; b < 65536; b += blockDim.x) hist[b] = 0;      // ready for the next pass
}

// candidates with key <= the k-th smallest key are vaccinated tomorrow
__global__ void k7_mark_kernel(DevCtx c, int64_t n_locals, const SelState *S, int32_t *vacc_day, uint8_t *selected, int *n_sel)
{
    const long long k = S->k_eff, nc = S->n_candidates;
    const unsigned long long thr = S->prefix;
    int cnt = 0;
    for (int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; a < n_locals; a += (int64_t)gridDim.x * blockDim.x) {
        bool sel = false;
        if (k > 0 && vacc_candidate(c, a)) {
            uint32_t w1;
            const unsigned long long key = vacc_key(c, a, &w1);
            sel = (k >= nc) || key <= thr;
            if (sel) {
                vacc_day[2 * a] = c.day.day + 1;                              // epidemiology.py:965
                vacc_day[2 * a + 1] = 42 + (int)__umulhi(w1, 79u);            // np.random.choice(arange(42, 121))
            }
        }
        selected[a] = sel;
        cnt += sel;
    }
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(FULLMASK, cnt, o);
    if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(n_sel, cnt);
}

// replay: the reference's own sample (agent ids + timesteps)
__global__ void k7_replay_kernel(DevCtx c, int64_t n, const int32_t *ids, const int32_t *ts, int32_t *vacc_day, uint8_t *selected, int *n_sel)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int a = ids[i];
        if (a < 0 || a >= c.n_agents) continue;
        vacc_day[2 * (int64_t)a] = c.day.day + 1; vacc_day[2 * (int64_t)a + 1] = ts[i];
        selected[a] = 1;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) *n_sel = (int)n;
}

__global__ void k7_dose_kernel(DevCtx c, int day, int64_t n_locals, int32_t *vacc_day)
{
    for (int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; a < n_locals; a += (int64_t)gridDim.x * blockDim.x) {
        const int vd = vacc_day[2 * a];
        if (vd == MTD_NONE) continue;
        if (day == vd) c.last_dose_day[a] = day;                                  // agents_vaccination_doses[a].append(day)
        else if (day > vd) { vacc_day[2 * a] = MTD_NONE; vacc_day[2 * a + 1] = MTD_NONE; }
    }
}

}  // namespace

// scratch: SelState (32 B) | n_sel int | pad ... ; hist: 65536 uint32
int launch_k7_schedule(const DevCtx &c, int64_t n_locals, long long k, int32_t *vacc_day, uint8_t *selected, void *scratch,
                       unsigned int *hist, cudaStream_t st, int sm_count)
{
    SelState *S = (SelState *)scratch;
    int *n_sel = (int *)((char *)scratch + 64);
    cudaMemsetAsync(scratch, 0, 128, st);
    cudaMemsetAsync(hist, 0, 65536 * sizeof(unsigned int), st);
    const int grid = sm_count * 8;
    k7_count_kernel<<<grid, 256, 0, st>>>(c, n_locals, S);
    k7_init_kernel<<<1, 1, 0, st>>>(S, k);
    for (int pass = 0; pass < 4; pass++) {
        k7_hist_kernel<<<grid, 256, 0, st>>>(c, n_locals, S, pass, hist);
        k7_pick_kernel<<<1, 256, 0, st>>>(S, pass, hist);
    }
    k7_mark_kernel<<<grid, 256, 0, st>>>(c, n_locals, S, vacc_day, selected, n_sel);
    return 11;
}

int launch_k7_replay(const DevCtx &c, int64_t n, const int32_t *ids, const int32_t *ts, int32_t *vacc_day, uint8_t *selected,
                     void *scratch, cudaStream_t st)
{
    int *n_sel = (int *)((char *)scratch + 64);
    cudaMemsetAsync(scratch, 0, 128, st);
    k7_replay_kernel<<<64, 256, 0, st>>>(c, n, ids, ts, vacc_day, selected, n_sel);
    return 1;
}

int launch_k7_dose(const DevCtx &c, int day, int64_t n_locals, int32_t *vacc_day, cudaStream_t st, int sm_count)
{
    k7_dose_kernel<<<sm_count * 4, 256, 0, st>>>(c, day, n_locals, vacc_day);
    return 1;
}
