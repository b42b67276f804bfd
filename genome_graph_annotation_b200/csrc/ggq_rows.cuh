// K1 (v3): batched BRWT::get_row (annotation/hierarchical_annotation/BRWT.cpp:26-53) for sm_100a.
//
// The engine first orders the batch by row id (radix sort of (row, query index) pairs): rank1 is monotone, so
// sorted rows walk EVERY node vector front to back — TLB- and DRAM-page-local instead of 33 random pages per
// row — and results are written back under the caller's query index.
//
// One warp owns a tile of `rb` consecutive (sorted) rows and a private LIFO stack of work items
// (t, group, r) = "probe the sibling group `group` at index r for row t of the tile".  Per iteration the warp
// pops up to 16 items; 2 lanes serve one item, each reading with ONE 256-bit load the 32-byte sector that
// holds the slots of two children (one coalesced 64-byte block per item).  A set bit of a leaf child marks the
// column's DFS rank in the row's bitmap (shared memory) and bumps the row's counter; a set bit of an internal
// child pushes the child's first group with r' = rank1(r) - 1.  When the stack drains, the bitmaps of the tile
// are scanned once in (row, word) order — which IS the reference's output order (DFS over child_nodes_,
// BRWT.cpp:45-50) — through per-row summary words (one bit per non-zero bitmap word), so that only the ~8 words
// per row that hold an id are visited instead of all 83.
//
// v3 vs v2 (profiles/r01: the kernel is instruction-issue bound, 27 warp instructions per probe): the two
// children of a lane are handled without divergent branches (rank computed unconditionally, leaf marks
// predicated), all shared-memory traffic uses precomputed 32-bit shared addresses, per-row counts come from
// shared-memory counters bumped during the descent (no counting pass), the bitmaps are cleared while they are
// read (no zeroing pass), and the emit works on non-zero words only instead of on every (row, 32-word chunk).
#pragma once
#include "ggq_device.cuh"

namespace ggq {

constexpr int kRowsWarps = 8;          // warps per CTA of the descent kernel
// Work items a warp probes per iteration (2 lanes each, 2 children per lane).  A variant with two items per lane pair
// and iteration (two independent slot loads in flight per lane, 64 registers -> 4 CTAs/SM instead of 6) was measured
// slower on B200: 12.3 vs 11.8 ms on the C4 shape at 1e9 rows.
constexpr int kItemsPerIter = 16;
constexpr int kRowCounters = 16;       // per-warp row counters (>= kMaxRowsPerWarp)
constexpr int kStatCounters = 4;       // per-warp kernel statistics: code probes, code escapes (flushed at kernel end)
constexpr u32 kCodeEscape = 255;       // leaf code "more than one leaf set below: descend"
constexpr u32 kCodedFlag = 1u << 30;   // wide group records: internal child with leaf codes

struct RowsArgs {
    const u64 *rows;          // query row ids in PROCESSING order (sorted by the engine, or the caller's order)
    const u32 *perm;          // processing position -> caller's query index (null: identity)
    u64 n;                    // number of queries
    u32 *tmp_cols;            // unordered chunk arena
    u64 tmp_cap;
    u64 *cursor;              // [0] chunk cursor, [1] tile ticket, [2] probe counter, [3] (node,row) touches, [4] overflow
                              // flag, [5] leaf-code probes, [6] leaf-code escapes
    u64 *row_tmp_off;         // per query (caller's index): start in tmp_cols
    u32 *row_cnt;             // per query (caller's index): number of columns
    u64 *spill;               // per-warp stack overflow area (cap_spill entries each)
    u32 cap_s, cap_spill;     // shared / global stack capacity per warp
    u32 bm_words;             // bitmap words per row
    u32 sum_words;            // summary words per row = ceil(bm_words / 32)
    u32 sum_magic;            // floor(2^32 / sum_words) + 1: e / sum_words by one mulhi
    u32 rb;                   // rows per warp tile
    // root fast path: the root is a group of one vector; when it is internal the seeding step probes it directly
    // (one 16-byte slot per row, all lanes busy) instead of spending a whole stack iteration on rb items
    u32 root_child;           // first group of the root's children, or kNoRootFast
    u32 root_sl;              // stride_log2 of group 0
    u32 root_multi_sb;        // root vector spans several superblocks
    u32 root_all_ones;        // the root vector has no zero: the seeding step needs no probe
    u64 root_base;            // slot base of group 0
    u32 any_more;             // some node has more than 4 children (continuation groups exist)
    int *err;
};
constexpr u32 kNoRootFast = 0xFFFFFFFFu;

// Dynamic shared memory of one CTA: [stacks: warps x cap_s u64][per warp: rb*bm_words bitmap words + rb*sum_words
// summary words + 16 row counters + 4 statistics counters (u32), padded to 16 bytes]
inline u32 rows_sum_words(u32 bm_words) { return (bm_words + 31) / 32; }
inline u32 rows_warp_words(u32 rb, u32 bm_words) {
    return (rb * (bm_words + rows_sum_words(bm_words)) + kRowCounters + kStatCounters + 3u) & ~3u;
}
inline size_t rows_smem_bytes(u32 cap_s, u32 rb, u32 bm_words) {
    return size_t(kRowsWarps) * (size_t(cap_s) * 8 + size_t(rows_warp_words(rb, bm_words)) * 4);
}

// Shared memory through 32-bit shared-window addresses (computed once per kernel).
__device__ __forceinline__ u64 lds64(u32 a) {
    u64 v;
    asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts64(u32 a, u64 v) { asm volatile("st.shared.u64 [%0], %1;" ::"r"(a), "l"(v) : "memory"); }
__device__ __forceinline__ u32 lds32(u32 a) {
    u32 v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts32(u32 a, u32 v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void reds_or(u32 a, u32 v) { asm volatile("red.shared.or.b32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void reds_add(u32 a, u32 v) { asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }

template <bool SPILL>
__device__ __forceinline__ u64 stk_get(u32 stk, const u64 *spill, u32 cap_s, u32 i) {
    if (!SPILL) return lds64(stk + i * 8);
    return i < cap_s ? lds64(stk + i * 8) : spill[i - cap_s];
}
template <bool SPILL>
__device__ __forceinline__ void stk_set(u32 stk, u64 *spill, u32 cap_s, u32 cap_total, u32 i, u64 v, int *err) {
    if (!SPILL) { sts64(stk + i * 8, v); return; }
    if (i < cap_s) sts64(stk + i * 8, v);
    else if (i < cap_total) spill[i - cap_s] = v;
    else *err = 1;
}

// ones strictly before position `bit` of the slot, given the word holding it (wsel = bit / 32, below = mask of
// the lower positions inside that word), plus the slot's rank header
__device__ __forceinline__ u32 slot_ones_before(const uint4 &s, u32 w, u32 wsel, u32 below) {
    u32 c = s.x + __popc(w & below);
    if (wsel >= 1) c += __popc(s.y);
    if (wsel >= 2) c += __popc(s.z);
    return c;
}

// Set bit of a leaf child: mark the DFS ranks of its columns in the row's bitmap, bump the row's counter.
// Besides the bitmap bit, a mark sets bit (word index & 31) of the row's summary word (word index >> 5): the emit
// visits only the bitmap words that are not zero.
template <bool MULTI_LEAF>
__device__ __forceinline__ void mark_leaf(const Img &img, u32 rank, u32 li, u32 bm_row, u32 sum_row, u32 cnt_addr) {
    if (!MULTI_LEAF) {
        reds_or(bm_row + (rank >> 5) * 4, 1u << (rank & 31));
        reds_or(sum_row + (rank >> 10) * 4, 1u << ((rank >> 5) & 31));
        reds_add(cnt_addr, 1u);
    } else {
        const u32 ncol = __ldg(img.lane_aux + li);
        for (u32 k = 0; k < ncol; ++k, ++rank) {
            reds_or(bm_row + (rank >> 5) * 4, 1u << (rank & 31));
            reds_or(sum_row + (rank >> 10) * 4, 1u << ((rank >> 5) & 31));
        }
        reds_add(cnt_addr, ncol);
    }
}

// Group record seen by one lane: where the group's blocks are, its shape flags (stride_log2 | cnt << 2 | more << 5 |
// multi-superblock << 6) and the lane's two children.  A child is `leaf bit | index`: for an internal child the
// index of its first group, for a leaf the DFS rank of its first column; all ones = held by another shard.
//   wide    (32 bytes per group, two 16-byte halves): 32-bit children, leaf bit 31      (Img::group_rec)
//   compact (16 bytes per group): 21-bit children, leaf bit 20 — trees with < 2^20 - 1 columns.  Half the table:
//            44 -> 22 KB for the C4 tree, which is what lets it stay in the ~28 KB of L1 left beside 6 CTAs' shared memory
template <bool COMPACT> struct RecFmt;
template <> struct RecFmt<false> { static constexpr u32 kLeaf = kLeafFlag, kRemote = kRemoteLane; };
template <> struct RecFmt<true> { static constexpr u32 kLeaf = 1u << 20, kRemote = (1u << 21) - 1u; };

template <bool COMPACT>
__device__ __forceinline__ void load_rec(const Img &img, u32 g, u32 half, u64 &base, u32 &flags, u32 &child0, u32 &child1,
                                         u32 &coded) {
    if (!COMPACT) {
        const uint4 rec = __ldg(img.group_rec + 2 * g + half);
        base = (u64(rec.y & 0xffu) << 32) | rec.x;
        flags = (rec.y >> (kGhStrideShift - 32)) & 0x7fu;
        // bit 30 of an internal child: the child carries leaf codes (a leaf / remote lane has bit 31 set)
        coded = ((rec.z >> 31) ? 0u : (rec.z >> 30) & 1u) | ((rec.w >> 31) ? 0u : (rec.w >> 29) & 2u);
        child0 = (rec.z >> 31) ? rec.z : rec.z & ~kCodedFlag;
        child1 = (rec.w >> 31) ? rec.w : rec.w & ~kCodedFlag;
    } else {
        // x: base / 4; y: flags | c0 << 7 | coded mask << 28; z: c1 | c2 << 21; w: c2 >> 11 | c3 << 10
        const uint4 rec = __ldg(img.group_rec16 + g);
        base = u64(rec.x) << 2;
        flags = rec.y & 0x7fu;
        const u32 m = (1u << 21) - 1u;
        child0 = half ? ((rec.z >> 21) | (rec.w << 11)) & m : (rec.y >> 7) & m;
        child1 = half ? (rec.w >> 10) & m : rec.z & m;
        coded = (rec.y >> (28 + 2 * half)) & 3u;
    }
}

// Leaf codes: an internal child Z whose subtree has at most 255 leaves can carry one byte per
// position p of its children's vectors: the DFS index (within the subtree) of the ONLY leaf set below Z at p, or 255
// when several are — then the subtree is descended as usual.  One 1-byte probe replaces the whole sub-descent
// (for the C4 tree: 3 levels, 22.9 of 33.4 probes per row) whenever a row has a single label under Z.
__device__ __forceinline__ u32 ld_code(const unsigned char *p) {
    u32 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::64B.u8 %0, [%1];" : "=r"(v) : "l"(p));   // (a plain miss fetches 128 B)
    return v;
}

// WIDE: the capacity layout (32-byte slots of 224 positions, ggq_layout.hpp): a lane reads one 32-byte sector per
// child — twice the sectors of the default layout, where one sector holds two children.
template <bool SPILL, bool MULTI_LEAF, bool COMPACT, bool WIDE = false>
__global__ void __launch_bounds__(kRowsWarps * 32)
brwt_rows_kernel(Img img, RowsArgs a) {
    typedef RecFmt<COMPACT> Fmt;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const u32 lt_mask = (1u << lane) - 1u;
    const u32 bm_total = a.rb * a.bm_words;
    const u32 smem0 = u32(__cvta_generic_to_shared(smem_raw));
    const u32 sum_total = a.rb * a.sum_words;
    const u32 warp_words = (bm_total + sum_total + kRowCounters + kStatCounters + 3u) & ~3u;
    u32 stk = smem0 + warp * a.cap_s * 8;
    u32 bm = smem0 + kRowsWarps * a.cap_s * 8 + warp * warp_words * 4;
    u32 sum0 = bm + bm_total * 4;
    u32 cnt0 = sum0 + sum_total * 4;
    // keep the per-warp bases in registers: ptxas otherwise rematerialises them (S2R + IMAD chains) inside the
    // descent loop, which is instruction-issue bound; a shuffle is opaque to it
    stk = __shfl_sync(0xffffffffu, stk, lane);
    bm = __shfl_sync(0xffffffffu, bm, lane);
    sum0 = __shfl_sync(0xffffffffu, sum0, lane);
    cnt0 = __shfl_sync(0xffffffffu, cnt0, lane);
    const u32 bm_row_bytes = a.bm_words * 4, sum_row_bytes = a.sum_words * 4;
    const u32 stat0 = cnt0 + kRowCounters * 4;
    u64 *spill = SPILL ? a.spill + (size_t(blockIdx.x) * kRowsWarps + warp) * a.cap_spill : nullptr;
    const u32 cap_total = a.cap_s + a.cap_spill;
    u32 idx = lane >> 1, half = lane & 1;   // 2 lanes serve one item; lane `half` owns children 2*half, 2*half+1
    const bool short_rows = img.n_rows < (1ull << 37);   // r / 96 by one 32-bit mulhi
    u32 probes = 0;    // lane 0
    u32 touches = 0;   // per lane: sibling vectors bit-tested (= V_t of SURVEY.md 8d, summed)

    // bitmaps and counters start clean; every tile leaves them clean
    for (u32 i = lane; i < warp_words; i += 32) sts32(bm + i * 4, 0);
    __syncwarp();

    for (;;) {
        u64 tile = 0;
        if (lane == 0) tile = atomicAdd(a.cursor + 1, 1ull);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        const u64 row0 = tile * a.rb;
        if (row0 >= a.n) break;
        const u32 nr = u32(min(u64(a.rb), a.n - row0));

        // seed the stack with the root probes of the valid rows
        u64 r = 0;
        bool valid = false;
        if (lane < nr) { r = a.rows[row0 + lane]; valid = r < img.n_rows; }
        const u32 vmask = __ballot_sync(0xffffffffu, valid);
        u32 top;
        if (a.root_child != kNoRootFast) {
            bool push = false;
            u64 it = 0;
            if (valid && a.root_all_ones) {
                // every row has a label (the usual case of a real annotation): rank1(r) - 1 = r, no probe at all
                it = (r << kItemRShift) | (u64(lane) << kItemGroupBits) | a.root_child;
                push = true;
                touches += 1;
            } else if (valid) {
                u64 blk; u32 bit;
                SlotT<WIDE>::div(r, blk, bit);
                const SlotT<WIDE> s = SlotT<WIDE>::load(img.slots + a.root_base + (blk << a.root_sl));
                if (s.test(bit)) {
                    u64 ones = u64(s.header()) + s.rank_before(bit);
                    if (a.root_multi_sb) ones += img.sb_base[img.lane_sb_off[0] + (blk >> kSuperShift)];
                    it = (ones << kItemRShift) | (u64(lane) << kItemGroupBits) | a.root_child;
                    push = true;
                }
                touches += 1;
            }
            const u32 pm = __ballot_sync(0xffffffffu, push);
            if (push) sts64(stk + __popc(pm & lt_mask) * 8, it);
            top = __popc(pm);
            if (!a.root_all_ones) probes += __popc(vmask);
        } else {
            if (valid) sts64(stk + __popc(vmask & lt_mask) * 8, (r << kItemRShift) | (u64(lane) << kItemGroupBits));
            top = __popc(vmask);
        }
        __syncwarp();

        while (top > 0) {
            const u32 c = min(top, u32(kItemsPerIter));
            const bool active = idx < c;
            u64 item = 0;
            if (active) item = stk_get<SPILL>(stk, spill, a.cap_s, top - 1 - idx);
            __syncwarp();   // every pop happens before any push below
            bool push_cont = false, push0 = false, push1 = false;
            u64 item0 = 0, item1 = 0;
            if (active) {
                const u32 ilo = u32(item);
                const u32 g = ilo & (kMaxGroups - 1);
                const u32 tq = ilo & ((kMaxRowsPerWarp - 1) << kItemGroupBits);   // t, in place
                u64 gbase;
                u32 flags, child0, child1, coded;
                load_rec<COMPACT>(img, g, half, gbase, flags, child0, child1, coded);
                const u32 sl = flags & 3u;
                const u32 cnt = (flags >> 2) & 7u;
                push_cont = half == 0 && ((flags >> 5) & 1u);
                if (half == 0) touches += cnt;
                if (2 * half < cnt) {
                    u64 blk; u32 bit;
                    if (WIDE) {
                        const u32 r32 = u32(item >> kItemRShift), q = r32 / kWideSlotBits;   // (vectors < 2^32 positions)
                        blk = q;
                        bit = r32 - q * kWideSlotBits;
                    } else if (short_rows) {
                        const u32 q = __umulhi(u32(item >> (kItemRShift + 5)), 0xAAAAAAABu) >> 1;
                        blk = q;
                        bit = u32(item >> kItemRShift) - q * kSlotBits;
                    } else {
                        const u64 rr = item >> kItemRShift;
                        blk = rr / kSlotBits;
                        bit = u32(rr - blk * kSlotBits);
                    }
                    const bool leaf0 = child0 & Fmt::kLeaf, leaf1 = child1 & Fmt::kLeaf;
                    const u32 t = tq >> kItemGroupBits;
                    const u32 bm_row = bm + t * bm_row_bytes, sum_row = sum0 + t * sum_row_bytes;
                    bool set0, set1;
                    u64 ones0, ones1;   // r' = rank1(r) - 1 = ones strictly before r (BRWT.cpp:43-50)
                    if (WIDE) {
                        const uint4 *p0 = img.slots + gbase + (blk << sl) + 4 * half;   // child 2*half; two 16-byte units per child
                        const bool has1 = 2 * half + 1 < cnt;
                        const SlotT<true> a0 = SlotT<true>::load(p0);
                        const SlotT<true> a1 = SlotT<true>::load(has1 ? p0 + 2 : p0);
                        set0 = a0.test(bit);
                        set1 = has1 && a1.test(bit);
                        ones0 = a0.header() + a0.rank_before(bit);
                        ones1 = a1.header() + a1.rank_before(bit);
                    } else {
                        const u64 si = gbase + (blk << sl) + 2 * half;   // slot of child 2*half
                        uint4 s0, s1;
                        ld_slot2(img.slots + (si & ~u64(1)), s0, s1);
                        if (si & 1) s0 = s1;                            // stride-1 groups: odd blocks sit in the upper half
                        const u32 wsel = bit >> 5, sh = bit & 31, below = (1u << sh) - 1u;
                        const u32 w0 = wsel == 0 ? s0.y : (wsel == 1 ? s0.z : s0.w);
                        const u32 w1 = wsel == 0 ? s1.y : (wsel == 1 ? s1.z : s1.w);
                        set0 = (w0 >> sh) & 1u;
                        set1 = (2 * half + 1 < cnt) && ((w1 >> sh) & 1u);
                        ones0 = slot_ones_before(s0, w0, wsel, below);
                        ones1 = slot_ones_before(s1, w1, wsel, below);
                    }
                    if (set0 && leaf0 && child0 != Fmt::kRemote) mark_leaf<MULTI_LEAF>(img, child0 & (Fmt::kLeaf - 1u), g * 4 + 2 * half, bm_row, sum_row, cnt0 + t * 4);
                    if (set1 && leaf1 && child1 != Fmt::kRemote) mark_leaf<MULTI_LEAF>(img, child1 & (Fmt::kLeaf - 1u), g * 4 + 2 * half + 1, bm_row, sum_row, cnt0 + t * 4);
                    push0 = set0 && !leaf0;
                    push1 = set1 && !leaf1;
                    if ((flags >> 6) & 1u) {     // some lane of the group spans several superblocks
                        const u64 sb = blk >> kSuperShift;
                        if (push0) ones0 += img.sb_base[img.lane_sb_off[g * 4 + 2 * half] + sb];   // entry 0 of a lane is 0
                        if (push1) ones1 += img.sb_base[img.lane_sb_off[g * 4 + 2 * half + 1] + sb];
                    }
                    if (coded) {   // leaf codes: a single label below the child is marked right here, no push
                        const bool c0 = push0 && (coded & 1u), c1 = push1 && (coded & 2u);
                        if (c0 | c1) {
                            const u32 li = g * 4 + 2 * half;
                            u32 code0 = kCodeEscape, code1 = kCodeEscape, lr0 = 0, lr1 = 0;
                            if (c0) {
                                const uint4 cr = __ldg(img.code_rec + li);
                                code0 = ld_code(img.codes + ((u64(cr.y) << 32) | cr.x) + ones0);
                                lr0 = cr.z;
                            }
                            if (c1) {
                                const uint4 cr = __ldg(img.code_rec + li + 1);
                                code1 = ld_code(img.codes + ((u64(cr.y) << 32) | cr.x) + ones1);
                                lr1 = cr.z;
                            }
                            if (code0 != kCodeEscape) { mark_leaf<false>(img, lr0 + code0, 0, bm_row, sum_row, cnt0 + t * 4); push0 = false; }
                            if (code1 != kCodeEscape) { mark_leaf<false>(img, lr1 + code1, 0, bm_row, sum_row, cnt0 + t * 4); push1 = false; }
                            reds_add(stat0, u32(c0) + u32(c1));                                          // code probes
                            const u32 esc = u32(c0 && push0) + u32(c1 && push1);
                            if (esc) reds_add(stat0 + 4, esc);                                           // escapes
                        }
                    }
                    item0 = (ones0 << kItemRShift) | tq | child0;   // (only pushed for internal children: plain group index)
                    item1 = (ones1 << kItemRShift) | tq | child1;
                }
            }
            const u32 m0 = __ballot_sync(0xffffffffu, push0);
            const u32 m1 = __ballot_sync(0xffffffffu, push1);
            u32 pos = top - c;
            u32 cont_mask = 0;
            if (a.any_more) cont_mask = __ballot_sync(0xffffffffu, push_cont);   // uniform: trees of arity <= 4 have no continuation groups
            const u32 new_top = pos + __popc(cont_mask) + __popc(m0) + __popc(m1);
            // The capacity comes from a replay of this loop on the host (simulate_stack), which is an estimate, not a
            // bound: never write past it.  One uniform compare per iteration; the engine reruns the batch with the
            // spill instantiation and a provable capacity when the flag is set (Matrix::stack_safe).
            if (new_top > cap_total) {
                if (lane == 0) *a.err = 1;
                break;
            }
            if (a.any_more) {
                if (push_cont)
                    stk_set<SPILL>(stk, spill, a.cap_s, cap_total, pos + __popc(cont_mask & lt_mask), item + 1, a.err);
                pos += __popc(cont_mask);
            }
            if (push0) stk_set<SPILL>(stk, spill, a.cap_s, cap_total, pos + __popc(m0 & lt_mask), item0, a.err);
            pos += __popc(m0);
            if (push1) stk_set<SPILL>(stk, spill, a.cap_s, cap_total, pos + __popc(m1 & lt_mask), item1, a.err);
            top = new_top;
            probes += c;
            __syncwarp();
        }

        // ---- emit: per-row counts from the counters, one chunk allocation per tile, then the column ids in
        // (row, rank) order
        u32 my_cnt = 0;   // lane t holds the count of row t
        if (lane < kRowCounters) { my_cnt = lds32(cnt0 + lane * 4); sts32(cnt0 + lane * 4, 0); }
        const u32 total = __reduce_add_sync(0xffffffffu, my_cnt);
        u64 base = 0;
        if (lane == 0 && total) base = atomicAdd(a.cursor, u64(total));
        base = __shfl_sync(0xffffffffu, base, 0);
        u32 pre = my_cnt;   // exclusive prefix of my_cnt over lanes
        #pragma unroll
        for (int o = 1; o < kRowCounters; o <<= 1) {
            const u32 v = __shfl_up_sync(0xffffffffu, pre, o);
            if (lane >= o) pre += v;
        }
        pre -= my_cnt;
        if (lane < nr) {
            const u64 q = a.perm ? u64(a.perm[row0 + lane]) : row0 + lane;
            a.row_tmp_off[q] = base + pre;
            a.row_cnt[q] = my_cnt;
        }
        if (total) {
            // Lane e of a pass owns summary word e = (row t, word group wg): it walks the set bits of the summary —
            // the non-zero bitmap words wg*32 + b of row t — once to count its ids and once to store them.  Flat
            // order (t, wg, b, bit) IS (row, rank) order, the reference's output order (BRWT.cpp:45-50).  Summary and
            // bitmap words are cleared as they are read: the next tile starts clean.
            const bool fits = base + total <= a.tmp_cap;   // else: the engine grows the arena and reruns
            const u32 n_sum = nr * a.sum_words;
            u64 pos = base;
            for (u32 e0 = 0; e0 < n_sum; e0 += 32) {
                const u32 e = e0 + lane;
                u32 sm = 0;
                if (e < n_sum) sm = lds32(sum0 + e * 4);
                if (!__any_sync(0xffffffffu, sm != 0)) continue;
                if (sm) sts32(sum0 + e * 4, 0);
                const u32 t = a.sum_words == 1 ? e : __umulhi(e, a.sum_magic);
                const u32 wg = e - t * a.sum_words;
                const u32 words = bm + t * bm_row_bytes + wg * 128;   // bitmap word wg*32 of row t
                u32 c = 0;
                for (u32 x = sm; x; x &= x - 1) c += __popc(lds32(words + (__ffs(x) - 1) * 4));
                u32 inc = c;
                #pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const u32 v = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o) inc += v;
                }
                u64 p = pos + inc - c;
                for (u32 x = sm; x; x &= x - 1) {
                    const u32 b = __ffs(x) - 1;
                    u32 bits = lds32(words + b * 4);
                    sts32(words + b * 4, 0);
                    const u32 *col = img.leaf_col + (wg * 32 + b) * 32;
                    while (bits) {
                        const u32 bit = __ffs(bits) - 1;
                        bits &= bits - 1;
                        if (fits) a.tmp_cols[p] = __ldg(col + bit);
                        ++p;
                    }
                }
                pos += __shfl_sync(0xffffffffu, inc, 31);
            }
        }
        __syncwarp();
    }
    if (lane == 0 && probes) atomicAdd(a.cursor + 2, u64(probes));
    touches = __reduce_add_sync(0xffffffffu, touches);
    if (lane == 0 && touches) atomicAdd(a.cursor + 3, u64(touches));
    if (lane < 2) {   // code probes -> cursor[5], escapes -> cursor[6]
        const u32 v = lds32(stat0 + lane * 4);
        if (v) atomicAdd(a.cursor + 5 + lane, u64(v));
    }
}

}  // namespace ggq
