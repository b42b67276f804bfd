// Device-side primitives on the slot layout (ggq_layout.hpp): bit test, rank, in-slot select.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "ggq_layout.hpp"

namespace ggq {

typedef unsigned long long u64;
typedef unsigned int u32;

// Everything a query kernel needs; passed by value.
struct Img {
    const uint4 *slots;
    const u64 *group_hot;
    const uint4 *group_rec;   // [g*2 + half] = { hot.lo, hot.hi, lane_info[4g+2*half], lane_info[4g+2*half+1] }
    const uint4 *group_rec16; // [g] = compact 16-byte record (ggq_rows.cuh: load_rec), null if the tree does not qualify
    const u64 *group_nblocks;
    const u32 *lane_info;
    const u32 *lane_aux;
    const u32 *lane_sb_off;
    const u64 *lane_sel_off;
    const u64 *lane_ones;
    const u64 *sb_base;
    const u32 *sel;
    const u32 *leaf_col;
    const u32 *col_path_off;
    const u32 *col_path;
    const unsigned char *codes;   // leaf codes of all coded lanes (ggq_rows.cuh), null if none
    const uint4 *code_rec;        // [g*4 + lane] = { codes offset lo, hi, DFS rank of the subtree's first leaf, leaves }
    u64 n_rows;
    u32 n_cols;
    u32 n_groups;
    u32 multi_leaf;
};

// Streaming 16-byte read of one slot: read-only path, do not allocate in L1 (the node tables live there).
__device__ __forceinline__ uint4 ld_slot(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// Streaming 32-byte read of two adjacent slots (one DRAM sector) — the 256-bit load of sm_100.
// `p` must be 32-byte aligned.  GGQ_L2_64B adds the .L2::64B qualifier: without it every L2 miss
// fetches 128 bytes from DRAM (measured: profiles/r01/mb_gather_ncu_dram_bytes.csv).
#ifndef GGQ_L2_64B
#define GGQ_L2_64B 1
#endif
__device__ __forceinline__ void ld_slot2(const uint4 *p, uint4 &a, uint4 &b) {
#if GGQ_L2_64B
    asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v8.u32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
#else
    asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
#endif
                 : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
                 : "l"(p));
}

// r / 96 and r % 96.  Fast path (r < 2^37, i.e. every vector shorter than 1.4e11 positions): one 32-bit mulhi.
__device__ __forceinline__ void div96(u64 r, u64 &blk, u32 &bit) {
    if (r < (1ull << 37)) {
        u32 q = __umulhi(u32(r >> 5), 0xAAAAAAABu) >> 1;
        blk = q;
        bit = u32(r) - q * kSlotBits;
    } else {
        blk = r / kSlotBits;
        bit = u32(r - blk * kSlotBits);
    }
}

// Cached variant for data with spatial reuse inside a warp (sorted select queries, streaming scans).
__device__ __forceinline__ uint4 ld_slot_cached(const uint4 *p) { return __ldg(p); }

__device__ __forceinline__ u32 slot_word(const uint4 &s, u32 bit) {
    return bit < 32 ? s.y : (bit < 64 ? s.z : s.w);
}

__device__ __forceinline__ bool slot_test(const uint4 &s, u32 bit) {
    return (slot_word(s, bit) >> (bit & 31)) & 1u;
}

// ones strictly before `bit` inside the slot (bit in [0, 96])
__device__ __forceinline__ u32 slot_rank_before(const uint4 &s, u32 bit) {
    u32 c = 0;
    if (bit >= 32) c += __popc(s.y);
    if (bit >= 64) c += __popc(s.z);
    if (bit >= 96) return c + __popc(s.w);
    u32 w = slot_word(s, bit);
    return c + __popc(w & ((1u << (bit & 31)) - 1u));
}

__device__ __forceinline__ u32 slot_popc(const uint4 &s) { return __popc(s.y) + __popc(s.z) + __popc(s.w); }

// position (0..95) of the k-th (0-based) set bit of the slot; k < slot_popc(s)
__device__ __forceinline__ u32 slot_select(const uint4 &s, u32 k) {
    u32 c0 = __popc(s.y), c1 = __popc(s.z);
    u32 w, base;
    if (k < c0) { w = s.y; base = 0; }
    else if (k < c0 + c1) { w = s.z; base = 32; k -= c0; }
    else { w = s.w; base = 64; k -= c0 + c1; }
    return base + __fns(w, 0, int(k) + 1);
}

__device__ __forceinline__ const uint4 *slot_addr(const Img &img, u64 hot, u64 blk, u32 lane) {
    u32 sl = u32(hot >> kGhStrideShift) & 3u;
    return img.slots + (hot & kGhBaseMask) + (blk << sl) + lane;
}

// absolute number of ones before block `blk` of lane li (= group*4+lane), given its slot
__device__ __forceinline__ u64 lane_rank_base(const Img &img, u32 li, u64 blk, const uint4 &s) {
    u64 r = s.x;
    u64 sb = blk >> kSuperShift;
    if (sb) r += img.sb_base[img.lane_sb_off[li] + sb];
    return r;
}

// splitmix64 finaliser — the portable generator of SURVEY.md §8(d)
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL;
    x ^= x >> 27; x *= 0x94d049bb133111ebULL;
    x ^= x >> 31;
    return x;
}

}  // namespace ggq
