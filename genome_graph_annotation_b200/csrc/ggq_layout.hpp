// Device image geometry shared by host code and kernels.  See DESIGN.md "Data layout in HBM".
//
// A *slot* is 16 bytes: { u32 rank_before, u32 bits[3] } = 96 positions of one bit vector plus the
// number of ones before them (relative to the vector's superblock base).  The children of one BRWT
// node all have the same length (popcount of the parent, BRWT_builders.cpp:51-66), and a query that
// descends from the parent probes ALL of them at the same index (BRWT.cpp:43-50) — so siblings are
// stored interleaved: a *group* holds up to 4 sibling vectors ("lanes"); block b of the group is
// `stride` consecutive slots (lane 0..cnt-1), i.e. one 16/32/64-byte naturally aligned unit, and ONE
// memory transaction answers bit-test + rank for every child of the node.
#pragma once
#include <cstdint>

namespace ggq {

constexpr uint32_t kSlotBits = 96;          // positions per slot
// Capacity layout ("wide slots"): a slot is 32 bytes = { u32 rank_before, u32 bits[7] } = 224
// positions — one DRAM sector per child and probe, 14.3 % header overhead instead of 33.3 %.  Blocks of a group are
// 32/64/128 bytes (the group's stride_log2 counts 16-byte units, so it is one more than in the default layout).
// Vectors are limited to 2^32 positions in this layout (one superblock: ones fit the 32-bit header).
constexpr uint32_t kWideSlotBits = 224;
template <bool WIDE> struct SlotGeom {
    static constexpr uint32_t kBits = WIDE ? kWideSlotBits : kSlotBits;   // positions per slot
    static constexpr uint32_t kWords = WIDE ? 7u : 3u;                     // data words per slot
    static constexpr uint32_t kQuads = WIDE ? 2u : 1u;                     // 16-byte units per slot
};
inline uint32_t slot_quads(uint32_t slot_bits) { return slot_bits == kWideSlotBits ? 2u : 1u; }
inline uint32_t slot_data_words(uint32_t slot_bits) { return slot_bits == kWideSlotBits ? 7u : 3u; }
constexpr uint32_t kGroupLanes = 4;         // sibling vectors interleaved per group
constexpr uint32_t kSuperShift = 25;        // superblock = 2^25 blocks = 96 * 2^25 positions (< 2^32 ones)
constexpr uint32_t kSelectShift = 10;       // one select sample per 1024 ones
constexpr uint32_t kLeafFlag = 0x80000000u;
constexpr uint32_t kRemoteLane = 0xFFFFFFFFu;  // lane_info of a subtree / leaf held by another shard: never descended
constexpr uint32_t kChunkBlocks = 256;      // blocks per scan chunk when rank headers are built

// group_hot[g] (u64): slot base (40 bits) | stride_log2 (2) | cnt (3) | more (1) | multi-superblock (1)
constexpr int kGhStrideShift = 40;
constexpr int kGhCntShift = 42;
constexpr int kGhMoreShift = 45;
constexpr int kGhSuperShift = 46;   // some lane of the group spans more than one superblock (rank needs sb_base)
constexpr uint64_t kGhBaseMask = (uint64_t(1) << 40) - 1;

inline uint64_t pack_group_hot(uint64_t base, uint32_t stride_log2, uint32_t cnt, uint32_t more) {
    return base | (uint64_t(stride_log2) << kGhStrideShift) | (uint64_t(cnt) << kGhCntShift) |
           (uint64_t(more) << kGhMoreShift);
}

// lane_info[g*4 + lane] (u32):
//   leaf:      kLeafFlag | rank of the leaf's first column in DFS (= reference get_row) order
//   internal:  index of the first group holding its children

// Work item of the descent kernel packed in 64 bits: r (40) | t (4) | group (20)
constexpr int kItemGroupBits = 20;
constexpr int kItemTBits = 4;
constexpr int kItemRShift = kItemGroupBits + kItemTBits;
constexpr uint64_t kMaxRows = uint64_t(1) << 40;
constexpr uint32_t kMaxGroups = 1u << kItemGroupBits;
constexpr uint32_t kMaxRowsPerWarp = 1u << kItemTBits;

}  // namespace ggq
