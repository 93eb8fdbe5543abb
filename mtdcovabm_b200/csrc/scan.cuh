// Device-wide exclusive scan (reduce / scan of block sums / apply) shared by K0 (CSR offsets) and K5 (compaction).
#pragma once
#include "common.cuh"

#define SCAN_TILE 2048
#define SCAN_THREADS 256

template <typename TIn>
__global__ void scan_reduce_kernel(const TIn *in, int64_t n_max, const int *n_dev, long long *block_sums)
{
    int64_t n = n_dev ? min((int64_t)*n_dev, n_max) : n_max;
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
    long long s = 0;
    for (int k = threadIdx.x; k < SCAN_TILE; k += SCAN_THREADS) {
        int64_t i = base + k;
        if (i < n) s += (long long)in[i];
    }
    __shared__ long long sh[SCAN_THREADS / 32];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(FULLMASK, s, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        long long t = 0;
        for (int w = 0; w < SCAN_THREADS / 32; w++) t += sh[w];
        block_sums[blockIdx.x] = t;
    }
}

static __global__ void scan_sums_kernel(long long *block_sums, int64_t n_blocks)
{
    // single block: exclusive scan of block_sums in place
    __shared__ long long sh[SCAN_THREADS];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t base = 0; base < n_blocks; base += SCAN_THREADS) {
        int64_t i = base + threadIdx.x;
        long long v = (i < n_blocks) ? block_sums[i] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < SCAN_THREADS; o <<= 1) {
            long long t = (threadIdx.x >= o) ? sh[threadIdx.x - o] : 0;
            __syncthreads();
            sh[threadIdx.x] += t;
            __syncthreads();
        }
        long long incl = sh[threadIdx.x];
        if (i < n_blocks) block_sums[i] = carry + incl - v;
        __syncthreads();
        if (threadIdx.x == SCAN_THREADS - 1) carry += incl;
        __syncthreads();
    }
}

template <typename TIn, typename TOut>
__global__ void scan_apply_kernel(const TIn *in, TOut *out, int64_t n_max, const int *n_dev, const long long *block_sums)
{
    // exclusive scan; out[n] receives the total.  Elements past n are not written.
    int64_t n = n_dev ? min((int64_t)*n_dev, n_max) : n_max;
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
    if (base > n) return;
    constexpr int PER = SCAN_TILE / SCAN_THREADS;
    long long v[PER];
    long long s = 0;
    int64_t i0 = base + (int64_t)threadIdx.x * PER;
#pragma unroll
    for (int k = 0; k < PER; k++) { v[k] = (i0 + k < n) ? (long long)in[i0 + k] : 0; s += v[k]; }
    // block exclusive scan of s
    __shared__ long long sh[SCAN_THREADS / 32];
    long long incl = s;
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 1; o < 32; o <<= 1) { long long t = __shfl_up_sync(FULLMASK, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) sh[w] = incl;
    __syncthreads();
    long long woff = 0;
    for (int k = 0; k < w; k++) woff += sh[k];
    long long run = block_sums[blockIdx.x] + woff + incl - s;
#pragma unroll
    for (int k = 0; k < PER; k++) {
        if (i0 + k <= n) out[i0 + k] = (TOut)run;   // element n gets the grand total
        run += v[k];
    }
}

template <typename TIn, typename TOut>
static inline int device_scan(const TIn *in, TOut *out, int64_t n_max, const int *n_dev, long long *block_sums, cudaStream_t st)
{
    int64_t nb = (n_max + 1 + SCAN_TILE - 1) / SCAN_TILE;   // +1: the total lands at out[n]
    scan_reduce_kernel<TIn><<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, n_max, n_dev, block_sums);
    scan_sums_kernel<<<1, SCAN_THREADS, 0, st>>>(block_sums, nb);
    scan_apply_kernel<TIn, TOut><<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, out, n_max, n_dev, block_sums);
    return 3;
}

