// mtd_occupancy_noalloc: the occupancy count of k3_transmission.cu (agents per SEIR state, dynamicstatistics.py:169-193)
// for callers that may not allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  No counterpart in the reference.
//
// The handle's own path (mtd_occupancy) accumulates into six counters that live in a buffer mtd_create allocated.
// This variant keeps its scratch -- one row of per-block partial counts per block plus an arrival ticket -- in a
// file-scope __device__ array of fixed size, so the entry point calls no cudaMalloc / cudaFree / allocating library
// routine at all.  Every GPU that loads the library holds the array (MTD_NOALLOC_SCRATCH_BYTES, 64.25 KB) until exit.
//
// Sharing rules: every call on a device uses the SAME array, and the array keeps the previous call's contents.  The
// partial rows a call reads are all rewritten by that call before they are read; the ticket must start at zero and is
// cleared on the caller's stream ahead of the kernel (and again by the kernel's last block).  Two calls that overlap
// on one GPU (different streams, or two host threads) race on the array -- serialise them.
#include "common.cuh"
#include "../../include/mtd_b200.h"
#include <mutex>
#include <stdio.h>

namespace {

constexpr int NA_THREADS = 256;
constexpr int64_t NA_CHUNK = 65536;                                   // agents per block: 256 threads x 16 loads of 16 states
constexpr int NA_MAX_BLOCKS = (int)(MTD_NOALLOC_MAX_AGENTS / NA_CHUNK);   // 1024
constexpr int NA_ROW = 8;                                             // u64 words per partial row (6 used)
constexpr int NA_TICKET = NA_MAX_BLOCKS * NA_ROW;                     // first word behind the partial rows
constexpr size_t NA_WORDS = (size_t)NA_TICKET + 32;                   // + 256 bytes: ticket word and padding

static_assert(NA_WORDS * sizeof(unsigned long long) == MTD_NOALLOC_SCRATCH_BYTES, "header and scratch size disagree");

// In host code this name is only a stand-in for the symbol; the device address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned long long g_occ_scratch[NA_WORDS];

__global__ void __launch_bounds__(NA_THREADS)
occupancy_noalloc_kernel(const int8_t *__restrict__ seir, int64_t n, unsigned long long *__restrict__ scratch,
                         long long *__restrict__ out6)
{
    __shared__ unsigned int sh[NA_ROW];
    __shared__ bool last;
    if (threadIdx.x < NA_ROW) sh[threadIdx.x] = 0;
    __syncthreads();
    unsigned int cnt[6] = { 0, 0, 0, 0, 0, 0 };
    const int64_t lo = (int64_t)blockIdx.x * NA_CHUNK;
    const int64_t hi = (lo + NA_CHUNK < n) ? lo + NA_CHUNK : n;
    const int64_t nvec = (hi - lo) >> 4;                               // lo is a multiple of 16: the loads stay aligned
    if ((reinterpret_cast<uintptr_t>(seir) & 15) == 0) {
        const uint4 *s16 = reinterpret_cast<const uint4 *>(seir + lo);
        for (int64_t q = threadIdx.x; q < nvec; q += NA_THREADS) {
            const uint4 v = s16[q];
            const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const uint32_t st = (w[k >> 2] >> (8 * (k & 3))) & 0xffu;
#pragma unroll
                for (int j = 0; j < 6; j++) cnt[j] += (st == (uint32_t)j);
            }
        }
        for (int64_t a = lo + (nvec << 4) + threadIdx.x; a < hi; a += NA_THREADS) {
            const int st = seir[a];
#pragma unroll
            for (int j = 0; j < 6; j++) cnt[j] += (st == j);
        }
    } else {                                                           // caller's pointer is not 16-byte aligned
        for (int64_t a = lo + threadIdx.x; a < hi; a += NA_THREADS) {
            const int st = seir[a];
#pragma unroll
            for (int j = 0; j < 6; j++) cnt[j] += (st == j);
        }
    }
#pragma unroll
    for (int j = 0; j < 6; j++) {
        const unsigned int t = __reduce_add_sync(0xffffffffu, cnt[j]);
        if ((threadIdx.x & 31) == 0 && t) atomicAdd(&sh[j], t);
    }
    __syncthreads();
    // this block's row of the scratch: rewritten in full by every call, so nothing of an earlier call survives in it
    if (threadIdx.x < NA_ROW) scratch[(size_t)blockIdx.x * NA_ROW + threadIdx.x] = sh[threadIdx.x];
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = (atomicAdd(&scratch[NA_TICKET], 1ULL) == (unsigned long long)gridDim.x - 1);
    __syncthreads();
    if (!last) return;
    __threadfence();
    // the last block to arrive sums the rows in block order (integer sums: any order gives the same result)
    unsigned long long acc[6] = { 0, 0, 0, 0, 0, 0 };
    for (int b = threadIdx.x; b < (int)gridDim.x; b += NA_THREADS) {
#pragma unroll
        for (int j = 0; j < 6; j++) acc[j] += __ldcg(&scratch[(size_t)b * NA_ROW + j]);
    }
    __shared__ unsigned long long tot[6];
    if (threadIdx.x < 6) tot[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 6; j++) {
        unsigned long long v = acc[j];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(&tot[j], v);
    }
    __syncthreads();
    if (threadIdx.x < 6) out6[threadIdx.x] = (long long)tot[threadIdx.x];
    if (threadIdx.x == 0) scratch[NA_TICKET] = 0;                      // leave the ticket at zero for the next call
}

// device address of the scratch array, resolved once per device with that device current
constexpr int NA_MAX_DEVICES = 64;
std::mutex g_mu;
unsigned long long *g_scratch_ptr[NA_MAX_DEVICES] = {};

int fail(char *err, int64_t err_cap, int code, const char *fmt, long long a = 0, long long b = 0)
{
    if (err && err_cap > 0) snprintf(err, (size_t)err_cap, fmt, a, b);
    return code;
}

int fail_cuda(char *err, int64_t err_cap, const char *what, cudaError_t e)
{
    if (err && err_cap > 0) snprintf(err, (size_t)err_cap, "mtd_occupancy_noalloc: %s: %s", what, cudaGetErrorString(e));
    return MTD_E_CUDA;
}

}  // namespace

extern "C" int mtd_occupancy_noalloc(int32_t device, const int8_t *seir, int64_t n_agents, int64_t *out6, void *stream,
                                     char *err, int64_t err_cap)
{
    if (err && err_cap > 0) err[0] = 0;
    if (!seir || !out6 || n_agents < 0) return fail(err, err_cap, MTD_E_ARG, "mtd_occupancy_noalloc: null pointer or negative n_agents");
    if (device < 0 || device >= NA_MAX_DEVICES) return fail(err, err_cap, MTD_E_ARG, "mtd_occupancy_noalloc: device %lld out of range", device);
    if (n_agents > MTD_NOALLOC_MAX_AGENTS)
        return fail(err, err_cap, MTD_E_CAPACITY,
                    "mtd_occupancy_noalloc: n_agents = %lld exceeds MTD_NOALLOC_MAX_AGENTS = %lld (the static scratch holds "
                    "one row per 65536 agents); use mtd_occupancy", (long long)n_agents, (long long)MTD_NOALLOC_MAX_AGENTS);
    int cur = -1;
    cudaError_t e = cudaGetDevice(&cur);
    if (e == cudaSuccess && cur != device) e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail_cuda(err, err_cap, "cudaSetDevice", e);
    unsigned long long *scratch;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (!g_scratch_ptr[device]) {
            // set-up, once per device, with that device current: the symbol's address ON THIS DEVICE.  (Passing
            // g_occ_scratch itself from host code would hand the kernel the host-side stand-in.)
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_occ_scratch);
            if (e != cudaSuccess || !p) {
                if (cur >= 0 && cur != device) cudaSetDevice(cur);
                return e != cudaSuccess ? fail_cuda(err, err_cap, "cudaGetSymbolAddress", e)
                                        : fail(err, err_cap, MTD_E_CUDA, "mtd_occupancy_noalloc: cudaGetSymbolAddress returned null");
            }
            g_scratch_ptr[device] = (unsigned long long *)p;
        }
        scratch = g_scratch_ptr[device];
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = n_agents == 0 ? 1 : (int)((n_agents + NA_CHUNK - 1) / NA_CHUNK);
    // the ticket must start at zero whatever an earlier (possibly aborted) call left behind
    e = cudaMemsetAsync(scratch + NA_TICKET, 0, sizeof(unsigned long long), st);
    if (e == cudaSuccess) {
        occupancy_noalloc_kernel<<<grid, NA_THREADS, 0, st>>>(seir, n_agents, scratch, (long long *)out6);
        e = cudaGetLastError();
    }
    if (cur >= 0 && cur != device) cudaSetDevice(cur);
    if (e != cudaSuccess) return fail_cuda(err, err_cap, "launch", e);
    return MTD_OK;
}
