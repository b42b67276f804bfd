// Microbenchmark: random aligned gathers of 16/32/64/128-byte blocks from a large HBM buffer, under
// different cudaLimitMaxL2FetchGranularity settings.  Answers: what useful GB/s can a random 64-byte
// probe stream reach on B200, and does the L2 fetch granularity double the DRAM traffic?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mb_gather mb_gather.cu && ./mb_gather
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__host__ __device__ inline uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL;
    x ^= x >> 27; x *= 0x94d049bb133111ebULL;
    x ^= x >> 31;
    return x;
}

// LANES lanes cooperate on one block of LANES*16 bytes; each thread does `iters` independent probes (UNROLL in flight)
template <int LANES, int UNROLL, bool NOALLOC>
__global__ void gather(const uint4 *buf, uint64_t n_blocks, uint64_t iters, uint64_t *sink) {
    uint64_t tid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    uint64_t item = tid / LANES;
    uint32_t j = tid % LANES;
    uint32_t acc = 0;
    for (uint64_t it = 0; it < iters; it += UNROLL) {
        uint4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            uint64_t b = mix64(item * 0x9E3779B97F4A7C15ULL + it + u) % n_blocks;
            const uint4 *p = buf + b * LANES + j;
            if (NOALLOC)
                asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(p));
            else
                v[u] = *p;
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345678) sink[0] = acc;
}

template <int LANES, int UNROLL, bool NOALLOC>
int run(const uint4 *buf, uint64_t bytes, uint64_t *sink, const char *tag) {
    uint64_t n_blocks = bytes / (LANES * 16);
    int threads = 256, grid = 148 * 8 * 4;
    uint64_t iters = 128;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    gather<LANES, UNROLL, NOALLOC><<<grid, threads>>>(buf, n_blocks, iters, sink);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    gather<LANES, UNROLL, NOALLOC><<<grid, threads>>>(buf, n_blocks, iters, sink);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    double probes = double(grid) * threads / LANES * iters;
    printf("  %-28s block=%3dB unroll=%d : %8.3f ms  %7.2f Gprobes/s  useful %7.1f GB/s\n", tag, LANES * 16, UNROLL, ms,
           probes / ms / 1e6, probes * LANES * 16 / ms / 1e6);
    return 0;
}

int main() {
    uint64_t bytes = 128ull << 30;
    uint4 *buf;
    uint64_t *sink;
    CK(cudaMalloc(&buf, bytes));
    CK(cudaMalloc(&sink, 8));
    CK(cudaMemset(buf, 1, bytes));
    size_t lim = 0;
    CK(cudaDeviceGetLimit(&lim, cudaLimitMaxL2FetchGranularity));
    printf("default cudaLimitMaxL2FetchGranularity = %zu\n", lim);
    uint64_t sizes[] = {64ull << 20, 256ull << 20, 1ull << 30, 4ull << 30, 16ull << 30, 32ull << 30, 64ull << 30, 128ull << 30};
    for (int si = 0; si < 8; ++si) {
        if (sizes[si] > bytes) break;
        printf("buffer %.2f GB\n", sizes[si] / 1e9);
        if (run<1, 4, true>(buf, sizes[si], sink, "nc.no_allocate")) return 1;
        if (run<2, 4, true>(buf, sizes[si], sink, "nc.no_allocate")) return 1;
        if (run<4, 4, true>(buf, sizes[si], sink, "nc.no_allocate")) return 1;
        if (run<8, 4, true>(buf, sizes[si], sink, "nc.no_allocate")) return 1;
    }
    return 0;
}
