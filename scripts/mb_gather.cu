// Microbenchmark: random aligned gathers of 32/64/128-byte blocks from a large HBM buffer.
// Answers, for the descent kernel's probe stream (DESIGN.md "Random-gather ceiling"):
//   * what useful GB/s a random BLOCK-byte probe stream reaches on B200 at a given number of loads in flight,
//   * whether L2 fetches more than the requested sectors from DRAM (run under
//     `ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum` to see the bytes per variant), and whether the
//     `.L2::64B` load qualifier or cudaLimitMaxL2FetchGranularity changes that.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mb_gather mb_gather.cu && ./mb_gather [GiB]
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__host__ __device__ inline uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL;
    x ^= x >> 27; x *= 0x94d049bb133111ebULL;
    x ^= x >> 31;
    return x;
}

enum { H_NA = 0, H_NA64 = 1, H_PLAIN = 2, H_NA256 = 3 };

template <int HINT>
__device__ __forceinline__ uint32_t ld16(const uint4 *p) {
    uint4 v;
    if (HINT == H_NA)
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    else if (HINT == H_NA64)
        asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    else
        asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v.x ^ v.y ^ v.z ^ v.w;
}
__device__ __forceinline__ uint32_t ld32(const uint4 *p) {
    uint32_t a, b, c, d, e, f, g, h;
    asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(a), "=r"(b), "=r"(c), "=r"(d), "=r"(e), "=r"(f), "=r"(g), "=r"(h) : "l"(p));
    return a ^ b ^ c ^ d ^ e ^ f ^ g ^ h;
}

// LANES lanes cooperate on one block; each lane loads PER x (16 or 32) bytes.  BLOCK = LANES * PER * VEC bytes.
template <int LANES, int PER, int VEC, int UNROLL, int HINT>
__global__ void __launch_bounds__(256) gather(const uint4 *buf, uint64_t n_blocks, uint64_t iters, uint64_t *sink) {
    uint64_t tid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    uint64_t item = tid / LANES;
    uint32_t j = tid % LANES;
    constexpr int U4_PER_BLOCK = LANES * PER * VEC / 16;
    uint32_t acc = 0;
    for (uint64_t it = 0; it < iters; it += UNROLL) {
        uint32_t v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            uint64_t b = mix64(item * 0x9E3779B97F4A7C15ULL + it + u) & (n_blocks - 1);  // n_blocks is a power of two
            const uint4 *p = buf + b * U4_PER_BLOCK + j * (PER * VEC / 16);
            uint32_t x = 0;
#pragma unroll
            for (int k = 0; k < PER; ++k) x ^= (VEC == 32) ? ld32(p + 2 * k) : ld16<HINT>(p + k);
            v[u] = x;
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) acc += v[u];
    }
    if (acc == 0x12345678) sink[0] = acc;
}

template <int LANES, int PER, int VEC, int UNROLL, int HINT>
int run(const uint4 *buf, uint64_t bytes, uint64_t *sink, const char *tag, int ctas_per_sm) {
    constexpr int BLOCK = LANES * PER * VEC;
    uint64_t n_blocks = bytes / BLOCK;
    int threads = 256, grid = 148 * ctas_per_sm;
    uint64_t total_probes = 1ull << 28;   // fixed work per variant
    uint64_t iters = total_probes / (uint64_t(grid) * threads / LANES);
    iters = (iters + UNROLL - 1) / UNROLL * UNROLL;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    gather<LANES, PER, VEC, UNROLL, HINT><<<grid, threads>>>(buf, n_blocks, iters / 4 / UNROLL * UNROLL + UNROLL, sink);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    gather<LANES, PER, VEC, UNROLL, HINT><<<grid, threads>>>(buf, n_blocks, iters, sink);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    double probes = double(grid) * threads / LANES * iters;
    printf("  %-22s block=%3dB lanes=%d x%d x%2dB unroll=%d ctas/sm=%d : %8.3f ms %7.2f Gprobe/s useful %7.1f GB/s  in-flight/SM %5.0f\n",
           tag, BLOCK, LANES, PER, VEC, UNROLL, ctas_per_sm, ms, probes / ms / 1e6, probes * BLOCK / ms / 1e6,
           double(ctas_per_sm) * threads / LANES * UNROLL);
    return 0;
}

int main(int argc, char **argv) {
    uint64_t gib = argc > 1 ? strtoull(argv[1], nullptr, 10) : 64;
    int gran = argc > 2 ? atoi(argv[2]) : 0;
    uint64_t bytes = gib << 30;
    uint4 *buf;
    uint64_t *sink;
    size_t lim = 0;
    CK(cudaDeviceGetLimit(&lim, cudaLimitMaxL2FetchGranularity));
    printf("default cudaLimitMaxL2FetchGranularity = %zu\n", lim);
    if (gran) {
        CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, gran));
        CK(cudaDeviceGetLimit(&lim, cudaLimitMaxL2FetchGranularity));
        printf("set cudaLimitMaxL2FetchGranularity = %zu\n", lim);
    }
    CK(cudaMalloc(&buf, bytes));
    CK(cudaMalloc(&sink, 8));
    CK(cudaMemset(buf, 1, bytes));
    printf("buffer %.1f GB\n", bytes / 1e9);
    if (argc > 3) {   // quick: the K1-shaped probe (2 lanes x 256-bit) at footprints argv[3..] GiB within the buffer
        for (int i = 3; i < argc; ++i) {
            const uint64_t fp = strtoull(argv[i], nullptr, 10) << 30;
            if (fp > bytes) continue;
            printf(" footprint %8.2f GiB\n", fp / double(1ull << 30));
            if (run<2, 1, 32, 4, H_NA>(buf, fp, sink, "nc.na v8", 8)) return 1;
        }
        return 0;
    }
    // footprint sweep: is the probe rate a TLB / page-walk limit?
    for (uint64_t fp = 1ull << 26; fp <= bytes; fp <<= 2) {
        printf(" footprint %8.2f GiB\n", fp / double(1ull << 30));
        if (run<4, 1, 16, 4, H_NA>(buf, fp, sink, "nc.na", 8)) return 1;
        if (run<2, 1, 32, 4, H_NA>(buf, fp, sink, "nc.na v8", 8)) return 1;
        if (run<4, 1, 32, 4, H_NA>(buf, fp, sink, "nc.na v8", 8)) return 1;
    }
    // occupancy sweep at the full footprint (latency-bound knee)
    printf(" occupancy sweep\n");
    if (run<4, 1, 16, 1, H_NA>(buf, bytes, sink, "nc.na", 1)) return 1;
    if (run<4, 1, 16, 1, H_NA>(buf, bytes, sink, "nc.na", 2)) return 1;
    if (run<4, 1, 16, 1, H_NA>(buf, bytes, sink, "nc.na", 4)) return 1;
    if (run<4, 1, 16, 2, H_NA>(buf, bytes, sink, "nc.na", 1)) return 1;
    if (run<4, 1, 16, 4, H_NA>(buf, bytes, sink, "nc.na", 1)) return 1;
    if (run<4, 1, 16, 8, H_NA>(buf, bytes, sink, "nc.na", 1)) return 1;
    printf(" variants at the full footprint\n");
    if (run<4, 1, 16, 4, H_NA64>(buf, bytes, sink, "nc.na.L2::64B", 8)) return 1;
    if (run<4, 1, 16, 4, H_PLAIN>(buf, bytes, sink, "ld.global", 8)) return 1;
    if (run<2, 1, 32, 8, H_NA>(buf, bytes, sink, "nc.na v8", 8)) return 1;
    if (run<1, 2, 32, 4, H_NA>(buf, bytes, sink, "nc.na v8", 8)) return 1;
    if (run<1, 4, 16, 4, H_PLAIN>(buf, bytes, sink, "ld.global", 8)) return 1;
    if (run<2, 1, 16, 4, H_NA>(buf, bytes, sink, "nc.na", 8)) return 1;
    if (run<1, 1, 32, 4, H_NA>(buf, bytes, sink, "nc.na v8", 8)) return 1;
    if (run<8, 1, 16, 4, H_NA>(buf, bytes, sink, "nc.na", 8)) return 1;
    return 0;
}
