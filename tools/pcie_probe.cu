// tools/pcie_probe.cu -- how fast do results leave the device?  copy engine (cudaMemcpyAsync D2H) against SM stores into
// mapped pinned host memory (the export kernel's way), alone and beside an H2D upload of the next day's intervals.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/pcie_probe.cu -o gpurun_out/pcie_probe && gpurun_out/pcie_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void stream_out(uint4 *dst, const uint4 *src, int64_t n16)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = src[i];
}

// a latency/issue-bound stand-in for K1: every thread chases a dependent chain of integer work with sparse global loads
__global__ void busy(const uint32_t *src, uint32_t *sink, int iters, int64_t n)
{
    uint32_t x = threadIdx.x + blockIdx.x * blockDim.x;
    for (int i = 0; i < iters; i++) {
        x = x * 1664525u + 1013904223u;
        if ((i & 63) == 0) x ^= src[(x >> 4) % n];
    }
    if (x == 12345u) sink[0] = x;
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

int main()
{
    const size_t out_bytes = 235u << 20, in_bytes = 141u << 20;
    void *d_out, *d_in, *h_out, *h_in, *h_out_dev;
    CK(cudaMalloc(&d_out, out_bytes)); CK(cudaMalloc(&d_in, in_bytes));
    CK(cudaHostAlloc(&h_out, out_bytes, cudaHostAllocMapped)); CK(cudaHostAlloc(&h_in, in_bytes, cudaHostAllocDefault));
    CK(cudaHostGetDevicePointer(&h_out_dev, h_out, 0));
    cudaStream_t s_out, s_in;
    CK(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking));
    cudaEvent_t e0, e1, f0, f1;
    cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&f0); cudaEventCreate(&f1);
    const int blocks_list[] = { 8, 32, 148, 296, 592 };
    for (int mode = 0; mode < 2; mode++) {            // 0: D2H alone, 1: beside an H2D
        for (int variant = -1; variant < 5; variant++) {      // -1: copy engine, else kernel with blocks_list[variant]
            float best = 1e9f, best_in = 1e9f;
            for (int rep = 0; rep < 5; rep++) {
                CK(cudaDeviceSynchronize());
                if (mode) { cudaEventRecord(f0, s_in); CK(cudaMemcpyAsync(d_in, h_in, in_bytes, cudaMemcpyHostToDevice, s_in)); cudaEventRecord(f1, s_in); }
                cudaEventRecord(e0, s_out);
                if (variant < 0) CK(cudaMemcpyAsync(h_out, d_out, out_bytes, cudaMemcpyDeviceToHost, s_out));
                else stream_out<<<blocks_list[variant], 256, 0, s_out>>>((uint4 *)h_out_dev, (const uint4 *)d_out, (int64_t)(out_bytes >> 4));
                cudaEventRecord(e1, s_out);
                CK(cudaDeviceSynchronize());
                float ms = 0, msi = 0;
                cudaEventElapsedTime(&ms, e0, e1);
                if (mode) cudaEventElapsedTime(&msi, f0, f1);
                if (ms < best) best = ms;
                if (mode && msi < best_in) best_in = msi;
            }
            printf("%s %-22s D2H %6.2f ms %6.1f GB/s", mode ? "with-h2d" : "alone   ",
                   variant < 0 ? "copy-engine" : "sm-stores", best, out_bytes / best / 1e6);
            if (variant >= 0) printf(" (blocks %d)", blocks_list[variant]);
            if (mode) printf("   H2D %6.2f ms %6.1f GB/s", best_in, in_bytes / best_in / 1e6);
            printf("\n");
        }
    }
    // does a D2H transfer slow down a compute kernel that runs beside it (and vice versa)?
    {
        cudaStream_t s_k; CK(cudaStreamCreateWithFlags(&s_k, cudaStreamNonBlocking));
        cudaEvent_t k0, k1; cudaEventCreate(&k0); cudaEventCreate(&k1);
        const int64_t n = out_bytes / 4;
        for (int variant = -2; variant <= 3; variant += 1) {
            if (variant >= 0 && variant != 3) continue;        // -2: kernel alone, -1: beside copy engine, 3: beside sm-stores (296 blocks)
            float best_k = 1e9f, best_c = 1e9f;
            for (int rep = 0; rep < 3; rep++) {
                CK(cudaDeviceSynchronize());
                cudaEventRecord(k0, s_k);
                busy<<<148 * 8, 256, 0, s_k>>>((const uint32_t *)d_out, (uint32_t *)d_in, 60000, n);
                cudaEventRecord(k1, s_k);
                cudaEventRecord(e0, s_out);
                if (variant == -1) CK(cudaMemcpyAsync(h_out, d_out, out_bytes, cudaMemcpyDeviceToHost, s_out));
                if (variant == 3) stream_out<<<296, 256, 0, s_out>>>((uint4 *)h_out_dev, (const uint4 *)d_out, (int64_t)(out_bytes >> 4));
                cudaEventRecord(e1, s_out);
                CK(cudaDeviceSynchronize());
                float a = 0, b = 0; cudaEventElapsedTime(&a, k0, k1); cudaEventElapsedTime(&b, e0, e1);
                if (a < best_k) best_k = a;
                if (b < best_c) best_c = b;
            }
            printf("busy kernel %s: kernel %6.2f ms, transfer %6.2f ms\n",
                   variant == -2 ? "alone            " : (variant == -1 ? "beside copy-engine" : "beside sm-stores  "), best_k, best_c);
        }
    }
    return 0;
}
