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

// ---- the same primitives, generic over the two slot layouts (ggq_layout.hpp: SlotGeom) -----------------------------
// SlotT<false> wraps the 16-byte slot above; SlotT<true> is the 32-byte slot of the capacity layout: q0 = { rank, w0, w1,
// w2 }, q1 = { w3, w4, w5, w6 }, always 32-byte aligned (one DRAM sector, one 256-bit load).
template <bool WIDE> struct SlotT;

template <> struct SlotT<false> {
    typedef SlotGeom<false> G;
    uint4 q;
    __device__ __forceinline__ static SlotT load(const uint4 *p) { SlotT s; s.q = ld_slot(p); return s; }
    __device__ __forceinline__ static SlotT load_cached(const uint4 *p) { SlotT s; s.q = __ldg(p); return s; }
    __device__ __forceinline__ static SlotT load_plain(const uint4 *p) { SlotT s; s.q = *p; return s; }
    __device__ __forceinline__ void store_header(uint4 *p, u32 h) { q.x = h; *p = q; }
    __device__ __forceinline__ static void store_words(uint4 *p, const u32 *w) { *p = make_uint4(0, w[0], w[1], w[2]); }
    __device__ __forceinline__ u32 header() const { return q.x; }
    __device__ __forceinline__ u32 word(u32 i) const { return i == 0 ? q.y : (i == 1 ? q.z : q.w); }
    __device__ __forceinline__ bool test(u32 bit) const { return slot_test(q, bit); }
    __device__ __forceinline__ u32 rank_before(u32 bit) const { return slot_rank_before(q, bit); }
    __device__ __forceinline__ u32 popc() const { return slot_popc(q); }
    __device__ __forceinline__ u32 select(u32 k) const { return slot_select(q, k); }
    __device__ __forceinline__ static void div(u64 r, u64 &blk, u32 &bit) { div96(r, blk, bit); }
};

template <> struct SlotT<true> {
    typedef SlotGeom<true> G;
    uint4 q0, q1;
    __device__ __forceinline__ static SlotT load(const uint4 *p) { SlotT s; ld_slot2(p, s.q0, s.q1); return s; }
    __device__ __forceinline__ static SlotT load_cached(const uint4 *p) { SlotT s; s.q0 = __ldg(p); s.q1 = __ldg(p + 1); return s; }
    __device__ __forceinline__ static SlotT load_plain(const uint4 *p) { SlotT s; s.q0 = p[0]; s.q1 = p[1]; return s; }
    __device__ __forceinline__ void store_header(uint4 *p, u32 h) { q0.x = h; p[0] = q0; }
    __device__ __forceinline__ static void store_words(uint4 *p, const u32 *w) {
        p[0] = make_uint4(0, w[0], w[1], w[2]);
        p[1] = make_uint4(w[3], w[4], w[5], w[6]);
    }
    __device__ __forceinline__ u32 header() const { return q0.x; }
    __device__ __forceinline__ u32 word(u32 i) const {
        return i < 3 ? (i == 0 ? q0.y : (i == 1 ? q0.z : q0.w)) : (i == 3 ? q1.x : (i == 4 ? q1.y : (i == 5 ? q1.z : q1.w)));
    }
    __device__ __forceinline__ bool test(u32 bit) const { return (word(bit >> 5) >> (bit & 31)) & 1u; }
    // ones strictly before `bit` inside the slot (bit in [0, 224])
    __device__ __forceinline__ u32 rank_before(u32 bit) const {
        const u32 ws = bit >> 5;
        u32 c = 0;
        if (ws > 0) c += __popc(q0.y);
        if (ws > 1) c += __popc(q0.z);
        if (ws > 2) c += __popc(q0.w);
        if (ws > 3) c += __popc(q1.x);
        if (ws > 4) c += __popc(q1.y);
        if (ws > 5) c += __popc(q1.z);
        if (ws > 6) return c + __popc(q1.w);
        return c + __popc(word(ws) & ((1u << (bit & 31)) - 1u));
    }
    __device__ __forceinline__ u32 popc() const {
        return __popc(q0.y) + __popc(q0.z) + __popc(q0.w) + __popc(q1.x) + __popc(q1.y) + __popc(q1.z) + __popc(q1.w);
    }
    // position (0..223) of the k-th (0-based) set bit; k < popc()
    __device__ __forceinline__ u32 select(u32 k) const {
        u32 base = 0;
        #pragma unroll
        for (u32 i = 0; i < 6; ++i) {
            const u32 c = __popc(word(i));
            if (k < c) return base + __fns(word(i), 0, int(k) + 1);
            k -= c;
            base += 32;
        }
        return base + __fns(q1.w, 0, int(k) + 1);
    }
    // every vector of this layout is shorter than 2^32 positions (assign_geometry)
    __device__ __forceinline__ static void div(u64 r, u64 &blk, u32 &bit) {
        const u32 r32 = u32(r), q = r32 / kWideSlotBits;
        blk = q;
        bit = r32 - q * kWideSlotBits;
    }
};

template <bool WIDE>
__device__ __forceinline__ const uint4 *slot_addr_t(const Img &img, u64 hot, u64 blk, u32 lane) {
    u32 sl = u32(hot >> kGhStrideShift) & 3u;
    return img.slots + (hot & kGhBaseMask) + (blk << sl) + lane * SlotGeom<WIDE>::kQuads;
}
// (host-side view of the same address, for kernels that take the arena pointer)
template <bool WIDE>
__device__ __forceinline__ u64 slot_index_t(u64 hot, u64 blk, u32 lane) {
    u32 sl = u32(hot >> kGhStrideShift) & 3u;
    return (hot & kGhBaseMask) + (blk << sl) + lane * SlotGeom<WIDE>::kQuads;
}

// absolute number of ones before block `blk` of lane li, given the slot's header
__device__ __forceinline__ u64 lane_rank_base_h(const Img &img, u32 li, u64 blk, u32 header) {
    u64 r = header;
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
