// Hand-written sm_100a kernels of the query engine.  HBM-bound integer work: no tensor cores.
//
//   pack_plain_kernel / chunk_sum_kernel / lane_scan_kernel / header_kernel   build the slot arena
//   brwt_rows_kernel      K1  batched BRWT::get_row  (BRWT.cpp:26-53)
//   gather_rows_kernel        temp chunks -> ordered CSR
//   brwt_get_kernel       K2  BRWT::get              (BRWT.cpp:9-24)
//   lane_enumerate_kernel K3a leaf call_ones         (BRWT.cpp:65-71)
//   lane_select_kernel    K3b rows[i] = select1(rows[i]+1) at one ancestor (BRWT.cpp:81-83)
#pragma once
#include "ggq_device.cuh"

namespace ggq {

constexpr int kRowsWarps = 8;          // warps per CTA of the descent kernel
constexpr int kItemsPerIter = 16;      // work items a warp probes per iteration (2 lanes each, 2 children per lane)

// ------------------------------------------------------------------------------------------
// Arena construction
// ------------------------------------------------------------------------------------------
struct ChunkTable {
    const u32 *chunk_lane;    // g*4 + lane
    const u32 *chunk_first;   // first block of the chunk
    const u32 *lane_chunk_off;  // per lane li: index of its first chunk (size 4*G + 1)
    u64 n_chunks;
};

// One CTA per chunk (kChunkBlocks threads): copy 96 plain bits per block into the slot data words.
__global__ void __launch_bounds__(kChunkBlocks)
pack_plain_kernel(uint4 *slots, const u64 *group_hot, const u64 *group_nblocks, ChunkTable ct,
                  const u32 *plain, const u64 *lane_plain_off) {
    u64 chunk = blockIdx.x + u64(blockIdx.y) * gridDim.x;
    if (chunk >= ct.n_chunks) return;
    u32 li = ct.chunk_lane[chunk];
    u32 g = li >> 2;
    u64 b = u64(ct.chunk_first[chunk]) + threadIdx.x;
    if (b >= group_nblocks[g]) return;
    u64 hot = group_hot[g];
    u32 sl = u32(hot >> kGhStrideShift) & 3u;
    const u32 *src = plain + lane_plain_off[li] + 3 * b;
    uint4 s;
    s.x = 0; s.y = src[0]; s.z = src[1]; s.w = src[2];
    slots[(hot & kGhBaseMask) + (b << sl) + (li & 3)] = s;
}

__device__ __forceinline__ u32 block_reduce_sum_256(u32 v, u32 *sh) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    u32 t = 0;
    #pragma unroll
    for (int i = 0; i < kChunkBlocks / 32; ++i) t += sh[i];
    __syncthreads();
    return t;
}

// pass A: ones per chunk
__global__ void __launch_bounds__(kChunkBlocks)
chunk_sum_kernel(const uint4 *slots, const u64 *group_hot, const u64 *group_nblocks, ChunkTable ct,
                 u32 *chunk_sum) {
    __shared__ u32 sh[kChunkBlocks / 32];
    u64 chunk = blockIdx.x + u64(blockIdx.y) * gridDim.x;
    if (chunk >= ct.n_chunks) return;
    u32 li = ct.chunk_lane[chunk];
    u32 g = li >> 2;
    u64 b = u64(ct.chunk_first[chunk]) + threadIdx.x;
    u64 hot = group_hot[g];
    u32 sl = u32(hot >> kGhStrideShift) & 3u;
    u32 c = 0;
    if (b < group_nblocks[g]) {
        uint4 s = slots[(hot & kGhBaseMask) + (b << sl) + (li & 3)];
        c = slot_popc(s);
    }
    u32 tot = block_reduce_sum_256(c, sh);
    if (threadIdx.x == 0) chunk_sum[chunk] = tot;
}

// pass B: one CTA per lane: exclusive scan of its chunk sums; also total ones and superblock bases
__global__ void __launch_bounds__(256)
lane_scan_kernel(const u32 *lane_chunk_off, const u32 *chunk_sum, u64 *chunk_base, u64 *lane_ones,
                 const u32 *lane_sb_off, u64 *sb_base, u32 n_lanes) {
    __shared__ u64 sh[256];
    __shared__ u64 carry_s;
    u32 li = blockIdx.x;
    if (li >= n_lanes) return;
    u32 c0 = lane_chunk_off[li], c1 = lane_chunk_off[li + 1];
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const u32 chunks_per_sb = (1u << kSuperShift) / kChunkBlocks;
    for (u32 base = c0; base < c1; base += 256) {
        u32 c = base + threadIdx.x;
        u64 v = c < c1 ? chunk_sum[c] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < 256; o <<= 1) {  // Hillis-Steele inclusive scan
            u64 add = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
            __syncthreads();
            sh[threadIdx.x] += add;
            __syncthreads();
        }
        u64 carry = carry_s;
        u64 excl = carry + sh[threadIdx.x] - v;
        if (c < c1) {
            chunk_base[c] = excl;
            u32 local = c - c0;
            if (local % chunks_per_sb == 0) sb_base[lane_sb_off[li] + local / chunks_per_sb] = excl;
        }
        __syncthreads();
        if (threadIdx.x == 255) carry_s = carry + sh[255];
        __syncthreads();
    }
    if (threadIdx.x == 0) lane_ones[li] = carry_s;
}

// pass C: rank headers (relative to the superblock base) and select samples
__global__ void __launch_bounds__(kChunkBlocks)
header_kernel(uint4 *slots, const u64 *group_hot, const u64 *group_nblocks, ChunkTable ct,
              const u64 *chunk_base, const u32 *lane_info, const u64 *lane_sel_off, u32 *sel,
              int want_select) {
    __shared__ u32 sh[kChunkBlocks];
    u64 chunk = blockIdx.x + u64(blockIdx.y) * gridDim.x;
    if (chunk >= ct.n_chunks) return;
    u32 li = ct.chunk_lane[chunk];
    u32 g = li >> 2;
    u64 nb = group_nblocks[g];
    u64 b = u64(ct.chunk_first[chunk]) + threadIdx.x;
    u64 hot = group_hot[g];
    u32 sl = u32(hot >> kGhStrideShift) & 3u;
    uint4 *p = slots + (hot & kGhBaseMask) + (b << sl) + (li & 3);
    uint4 s = make_uint4(0, 0, 0, 0);
    u32 c = 0;
    if (b < nb) { s = *p; c = slot_popc(s); }
    sh[threadIdx.x] = c;
    __syncthreads();
    for (int o = 1; o < kChunkBlocks; o <<= 1) {
        u32 add = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
        __syncthreads();
        sh[threadIdx.x] += add;
        __syncthreads();
    }
    if (b >= nb) return;
    u64 abs_before = chunk_base[chunk] + sh[threadIdx.x] - c;
    // the superblock base is the chunk base of the first chunk of the superblock
    const u32 chunks_per_sb = (1u << kSuperShift) / kChunkBlocks;
    u32 local = u32(chunk - ct.lane_chunk_off[li]);
    u64 sbb = chunk_base[chunk - (local % chunks_per_sb)];
    s.x = u32(abs_before - sbb);
    *p = s;
    bool internal = !(lane_info[li] & kLeafFlag) || g == 0;
    if (want_select && internal) {
        u32 *my_sel = sel + lane_sel_off[li];
        u64 first = (abs_before + (1ull << kSelectShift) - 1) >> kSelectShift;
        if ((first << kSelectShift) < abs_before + c) my_sel[first] = u32(b);
        // (a 96-bit block holds at most one multiple of 1024)
        if (b == nb - 1) {  // sentinel sample after the last real one
            u64 ones = abs_before + c;
            my_sel[(ones + (1ull << kSelectShift) - 1) >> kSelectShift] = u32(b);
        }
    }
}

// ------------------------------------------------------------------------------------------
// K1: batched get_row.  The engine first orders the batch by row id (radix sort of (row, query index)
// pairs): rank1 is monotone, so sorted rows walk EVERY node vector front to back — TLB- and DRAM-page-local
// instead of 33 random pages per row — and results are written back under the caller's query index.
// One warp owns a batch of `rb` consecutive (sorted) rows and a private LIFO stack of work items
// (t, group, r) = "probe the sibling group `group` at index r for query t".  Per iteration the warp pops up
// to 16 items; 2 lanes serve one item, each reading with ONE 256-bit load the 32-byte sector that holds the
// slots of two children (one coalesced 64-byte block per item).  A set bit of a leaf child marks the column's
// DFS rank in the row's bitmap (shared memory); a set bit of an internal child pushes the child's first group
// with r' = rank1(r) - 1.  When the stack drains, the bitmaps are enumerated in rank order, which IS the
// reference's output order (DFS over child_nodes_, BRWT.cpp:45-50).
// ------------------------------------------------------------------------------------------
struct RowsArgs {
    const u64 *rows;          // query row ids in PROCESSING order (sorted by the engine, or the caller's order)
    const u32 *perm;          // processing position -> caller's query index (null: identity)
    u64 n;                    // number of queries
    u32 *tmp_cols;            // unordered chunk arena
    u64 tmp_cap;
    u64 *cursor;              // [0] chunk cursor, [1] batch ticket, [2] probe counter, [3] (node,row) touches
    u64 *row_tmp_off;         // per query (caller's index): start in tmp_cols
    u32 *row_cnt;             // per query (caller's index): number of columns
    u64 *spill;               // per-warp stack overflow area (cap_spill entries each)
    u32 cap_s, cap_spill;     // shared / global stack capacity per warp
    u32 bm_words;             // bitmap words per row
    u32 rb;                   // rows per warp batch
    int *err;
};

// Stack access.  SPILL = false: the shared-memory stack is provably large enough (simulate_stack), no checks.
template <bool SPILL>
__device__ __forceinline__ u64 stk_get(const u64 *s_stack, const u64 *spill, u32 cap_s, u32 i) {
    if (!SPILL) return s_stack[i];
    return i < cap_s ? s_stack[i] : spill[i - cap_s];
}
template <bool SPILL>
__device__ __forceinline__ void stk_set(u64 *s_stack, u64 *spill, u32 cap_s, u32 cap_total, u32 i, u64 v, int *err) {
    if (!SPILL) { s_stack[i] = v; return; }
    if (i < cap_s) s_stack[i] = v;
    else if (i < cap_total) spill[i - cap_s] = v;
    else *err = 1;
}

// One child of a probed group: bit test; a set leaf bit marks the row's bitmap, a set internal bit yields the
// work item of the child's first group with r' = rank1(r) - 1 (BRWT.cpp:43-50).
__device__ __forceinline__ bool visit_child(const Img &img, const uint4 &s, u32 bit, u64 blk, u32 li, u32 info,
                                            bool multi_sb, u32 t, u32 *s_bm_row, u64 *child_item) {
    if (!slot_test(s, bit)) return false;
    if (info & kLeafFlag) {
        if (info == kRemoteLane) return false;   // held by another shard (column sharding)
        u32 rank = info & ~kLeafFlag;
        if (!img.multi_leaf) {
            atomicOr(s_bm_row + (rank >> 5), 1u << (rank & 31));
        } else {
            u32 ncol = __ldg(img.lane_aux + li);
            for (u32 k = 0; k < ncol; ++k, ++rank) atomicOr(s_bm_row + (rank >> 5), 1u << (rank & 31));
        }
        return false;
    }
    u64 ones = u64(s.x) + slot_rank_before(s, bit);
    if (multi_sb) ones += img.sb_base[img.lane_sb_off[li] + (blk >> kSuperShift)];   // entry 0 of a lane is 0
    *child_item = (ones << kItemRShift) | (u64(t) << kItemGroupBits) | info;
    return true;
}

template <bool SPILL>
__global__ void __launch_bounds__(kRowsWarps * 32)
brwt_rows_kernel(Img img, RowsArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const u32 lt_mask = (1u << lane) - 1u;
    u64 *s_stack = reinterpret_cast<u64 *>(smem_raw) + size_t(warp) * a.cap_s;
    u32 *s_bm = reinterpret_cast<u32 *>(smem_raw + size_t(kRowsWarps) * a.cap_s * 8) + size_t(warp) * a.rb * a.bm_words;
    u64 *spill = SPILL ? a.spill + (size_t(blockIdx.x) * kRowsWarps + warp) * a.cap_spill : nullptr;
    const u32 cap_total = a.cap_s + a.cap_spill;
    const u32 bm_total = a.rb * a.bm_words;
    const u32 idx = lane >> 1, half = lane & 1;   // 2 lanes serve one item; lane `half` owns children 2*half, 2*half+1
    u32 probes = 0;    // lane 0
    u32 touches = 0;   // per lane: sibling vectors bit-tested (= V_t of SURVEY.md 8d, summed)

    for (;;) {
        u64 batch = 0;
        if (lane == 0) batch = atomicAdd(a.cursor + 1, 1ull);
        batch = __shfl_sync(0xffffffffu, batch, 0);
        const u64 row0 = batch * a.rb;
        if (row0 >= a.n) break;
        const u32 nr = u32(min(u64(a.rb), a.n - row0));

        for (u32 i = lane; i < bm_total; i += 32) s_bm[i] = 0;
        // seed the stack with the root probes of the valid rows
        u64 r = 0;
        bool valid = false;
        if (lane < nr) { r = a.rows[row0 + lane]; valid = r < img.n_rows; }
        const u32 vmask = __ballot_sync(0xffffffffu, valid);
        if (valid) s_stack[__popc(vmask & lt_mask)] = (r << kItemRShift) | (u64(lane) << kItemGroupBits);
        u32 top = __popc(vmask);
        __syncwarp();

        while (top > 0) {
            const u32 c = min(top, u32(kItemsPerIter));
            const bool active = idx < c;
            u64 item = 0;
            if (active) item = stk_get<SPILL>(s_stack, spill, a.cap_s, top - 1 - idx);
            __syncwarp();
            const u32 g = u32(item) & (kMaxGroups - 1);
            const u32 t = u32(item >> kItemGroupBits) & (kMaxRowsPerWarp - 1);
            const u64 rr = item >> kItemRShift;
            bool push_cont = false, push0 = false, push1 = false;
            u64 item0 = 0, item1 = 0;
            if (active) {
                const uint4 rec = __ldg(img.group_rec + 2 * g + half);
                const u32 sl = (rec.y >> (kGhStrideShift - 32)) & 3u;
                const u32 cnt = (rec.y >> (kGhCntShift - 32)) & 7u;
                push_cont = half == 0 && ((rec.y >> (kGhMoreShift - 32)) & 1u);
                const bool multi_sb = (rec.y >> (kGhSuperShift - 32)) & 1u;
                if (half == 0) touches += cnt;
                if (2 * half < cnt) {
                    u64 blk; u32 bit;
                    div96(rr, blk, bit);
                    const u64 base = (u64(rec.y & 0xffu) << 32) | rec.x;
                    const u64 si = base + (blk << sl) + 2 * half;   // slot of child 2*half
                    uint4 s0, s1;
                    ld_slot2(img.slots + (si & ~u64(1)), s0, s1);
                    if (si & 1) s0 = s1;                            // stride-1 groups: odd blocks sit in the upper half
                    u32 *bm_row = s_bm + t * a.bm_words;
                    push0 = visit_child(img, s0, bit, blk, g * 4 + 2 * half, rec.z, multi_sb, t, bm_row, &item0);
                    if (2 * half + 1 < cnt)
                        push1 = visit_child(img, s1, bit, blk, g * 4 + 2 * half + 1, rec.w, multi_sb, t, bm_row, &item1);
                }
            }
            const u32 cont_mask = __ballot_sync(0xffffffffu, push_cont);
            const u32 m0 = __ballot_sync(0xffffffffu, push0);
            const u32 m1 = __ballot_sync(0xffffffffu, push1);
            u32 pos = top - c;
            if (push_cont)
                stk_set<SPILL>(s_stack, spill, a.cap_s, cap_total, pos + __popc(cont_mask & lt_mask), item + 1, a.err);
            pos += __popc(cont_mask);
            if (push0) stk_set<SPILL>(s_stack, spill, a.cap_s, cap_total, pos + __popc(m0 & lt_mask), item0, a.err);
            pos += __popc(m0);
            if (push1) stk_set<SPILL>(s_stack, spill, a.cap_s, cap_total, pos + __popc(m1 & lt_mask), item1, a.err);
            top = pos + __popc(m1);
            if (SPILL && top > cap_total) top = cap_total;  // overflow already flagged in *err
            probes += c;
            __syncwarp();
        }

        // ---- emit: per-row counts, one chunk allocation per batch, then the column ids in rank order
        u32 my_cnt = 0;   // lane t holds the count of row t
        u32 total = 0;
        for (u32 t = 0; t < nr; ++t) {
            u32 c = 0;
            for (u32 w = lane; w < a.bm_words; w += 32) c += __popc(s_bm[t * a.bm_words + w]);
            c = __reduce_add_sync(0xffffffffu, c);
            if (lane == t) my_cnt = c;
            total += c;
        }
        u64 base = 0;
        if (lane == 0 && total) base = atomicAdd(a.cursor, u64(total));
        base = __shfl_sync(0xffffffffu, base, 0);
        // exclusive prefix of my_cnt over lanes
        u32 pre = my_cnt;
        for (int o = 1; o < 32; o <<= 1) {
            u32 v = __shfl_up_sync(0xffffffffu, pre, o);
            if (lane >= o) pre += v;
        }
        pre -= my_cnt;
        if (lane < nr) {
            const u64 q = a.perm ? u64(a.perm[row0 + lane]) : row0 + lane;
            a.row_tmp_off[q] = base + pre;
            a.row_cnt[q] = my_cnt;
        }
        const bool fits = base + total <= a.tmp_cap;
        if (fits && total) {
            for (u32 t = 0; t < nr; ++t) {
                u64 pos = base + __shfl_sync(0xffffffffu, pre, t);
                if (__shfl_sync(0xffffffffu, my_cnt, t) == 0) continue;
                for (u32 w0 = 0; w0 < a.bm_words; w0 += 32) {
                    u32 w = w0 + lane;
                    u32 bits = w < a.bm_words ? s_bm[t * a.bm_words + w] : 0u;
                    if (!__any_sync(0xffffffffu, bits != 0)) continue;
                    const u32 c = __popc(bits);
                    if (!__any_sync(0xffffffffu, c > 1)) {   // common case: at most one id per word
                        const u32 nz = __ballot_sync(0xffffffffu, c != 0);
                        if (c) a.tmp_cols[pos + __popc(nz & lt_mask)] = __ldg(img.leaf_col + w * 32 + __ffs(bits) - 1);
                        pos += __popc(nz);
                        continue;
                    }
                    u32 inc = c;
                    for (int o = 1; o < 32; o <<= 1) {
                        u32 v = __shfl_up_sync(0xffffffffu, inc, o);
                        if (lane >= o) inc += v;
                    }
                    u64 p = pos + inc - c;
                    while (bits) {
                        u32 b = __ffs(bits) - 1;
                        bits &= bits - 1;
                        a.tmp_cols[p++] = __ldg(img.leaf_col + w * 32 + b);
                    }
                    pos += __shfl_sync(0xffffffffu, inc, 31);
                }
            }
        }
        __syncwarp();
    }
    if (lane == 0 && probes) atomicAdd(a.cursor + 2, u64(probes));
    touches = __reduce_add_sync(0xffffffffu, touches);
    if (lane == 0 && touches) atomicAdd(a.cursor + 3, u64(touches));
}

// CSR assembly in the caller's query order: 8 lanes copy one row from its chunk in the arena.
template <typename OutT>
__global__ void gather_rows_kernel(const u32 *tmp_cols, const u64 *row_tmp_off, const u64 *offsets,
                                   u64 n, OutT *out) {
    const u64 q = (u64(blockIdx.x) * blockDim.x + threadIdx.x) >> 3;
    const u32 sub = threadIdx.x & 7;
    if (q >= n) return;
    const u64 dst0 = offsets[q], len = offsets[q + 1] - dst0;
    const u64 src0 = row_tmp_off[q];
    for (u64 i = sub; i < len; i += 8) out[dst0 + i] = OutT(tmp_cols[src0 + i]);
}

// iota for the (row id, query index) sort
__global__ void iota_kernel(u32 *p, u64 n) {
    u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) p[i] = u32(i);
}

// count_labels: histogram of column ids (StaticBinRelAnnotator::count_labels, annotate_static.cpp:149-162).
// Privatised in shared memory when the matrix has at most `smem_cols` columns, else straight global atomics.
__global__ void __launch_bounds__(512)
col_hist_kernel(const u32 *ids, u64 n_ids, u32 n_cols, u32 smem_cols, u64 *counts) {
    extern __shared__ u32 s_hist[];
    const bool priv = n_cols <= smem_cols;
    if (priv) {
        for (u32 i = threadIdx.x; i < n_cols; i += blockDim.x) s_hist[i] = 0;
        __syncthreads();
    }
    const u64 stride = u64(gridDim.x) * blockDim.x;
    for (u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x; i < n_ids; i += stride) {
        const u32 c = ids[i];
        if (priv) atomicAdd(s_hist + c, 1u);
        else atomicAdd(counts + c, 1ull);
    }
    if (priv) {
        __syncthreads();
        for (u32 i = threadIdx.x; i < n_cols; i += blockDim.x) {
            const u32 v = s_hist[i];
            if (v) atomicAdd(counts + i, u64(v));
        }
    }
}

// chunk-local CSR offsets -> global offsets (pipelined host API)
__global__ void add_base_kernel(u64 *p, u64 n, u64 base) {
    u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) p[i] += base;
}

// slice_rows: CSR -> rows separated by UINT64_MAX
__global__ void slice_rows_kernel(const u64 *offsets, const u32 *cols, u64 n, u64 *out) {
    u64 t = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= n) return;
    u64 b = offsets[t], e = offsets[t + 1];
    u64 o = b + t;
    for (u64 i = b; i < e; ++i) out[o++] = cols[i];
    out[o] = ~0ull;
}

// ------------------------------------------------------------------------------------------
// K2: get(row, column).  One thread per query walks the column's root-to-leaf lane path.
// ------------------------------------------------------------------------------------------
__global__ void brwt_get_kernel(Img img, const u64 *rows, const u64 *cols, u64 fixed_col, u64 n,
                                unsigned char *out) {
    u64 t = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= n) return;
    u64 r = rows[t];
    u64 col = cols ? cols[t] : fixed_col;
    bool res = false;
    if (r < img.n_rows && col < img.n_cols) {
        u32 p0 = img.col_path_off[col], p1 = img.col_path_off[col + 1];
        res = true;
        for (u32 p = p0; p < p1; ++p) {
            u32 li = img.col_path[p];
            u64 hot = __ldg(img.group_hot + (li >> 2));
            u64 blk = r / kSlotBits;
            u32 bit = u32(r - blk * kSlotBits);
            uint4 s = ld_slot(slot_addr(img, hot, blk, li & 3));
            if (!slot_test(s, bit)) { res = false; break; }
            if (p + 1 < p1) r = lane_rank_base(img, li, blk, s) + slot_rank_before(s, bit);
        }
    }
    out[t] = res;
}

// ------------------------------------------------------------------------------------------
// K3: get_column
// ------------------------------------------------------------------------------------------
// a: positions of the ones of one lane, ascending (one thread per block; the rank header is the output index)
__global__ void lane_enumerate_kernel(Img img, u32 li, u64 *out) {
    u32 g = li >> 2;
    u64 nb = img.group_nblocks[g];
    u64 b = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= nb) return;
    u64 hot = img.group_hot[g];
    uint4 s = ld_slot_cached(slot_addr(img, hot, b, li & 3));
    u64 o = lane_rank_base(img, li, b, s);
    u64 pos = b * kSlotBits;
    u32 w = s.y;
    while (w) { out[o++] = pos + __ffs(w) - 1; w &= w - 1; }
    w = s.z;
    while (w) { out[o++] = pos + 32 + __ffs(w) - 1; w &= w - 1; }
    w = s.w;
    while (w) { out[o++] = pos + 64 + __ffs(w) - 1; w &= w - 1; }
}

// select of the q-th (0-based) one of lane li
__device__ __forceinline__ u64 lane_select0(const Img &img, u32 li, u64 hot, u64 q) {
    const u32 *sel = img.sel + img.lane_sel_off[li];
    u64 si = q >> kSelectShift;
    u64 lo = sel[si], hi = sel[si + 1];
    // last block in [lo, hi] whose absolute header is <= q
    while (lo < hi) {
        u64 mid = (lo + hi + 1) >> 1;
        uint4 s = ld_slot_cached(slot_addr(img, hot, mid, li & 3));
        if (lane_rank_base(img, li, mid, s) <= q) lo = mid; else hi = mid - 1;
    }
    uint4 s = ld_slot_cached(slot_addr(img, hot, lo, li & 3));
    u64 before = lane_rank_base(img, li, lo, s);
    return lo * kSlotBits + slot_select(s, u32(q - before));
}

// b: rows[i] = select1(rows[i] + 1) in lane li  (BRWT.cpp:81-83)
__global__ void lane_select_kernel(Img img, u32 li, u64 *rows, u64 n) {
    u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    u64 hot = __ldg(img.group_hot + (li >> 2));
    rows[i] = lane_select0(img, li, hot, rows[i]);
}

}  // namespace ggq
