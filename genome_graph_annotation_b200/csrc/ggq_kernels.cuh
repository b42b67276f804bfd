// Hand-written sm_100a kernels of the query engine.  HBM-bound integer work: no tensor cores.
//
//   pack_plain_kernel / chunk_sum_kernel / lane_scan_kernel / header_kernel   build the slot arena
//   brwt_rows_kernel      K1  batched BRWT::get_row  (BRWT.cpp:26-53)  -> ggq_rows.cuh
//   gather_rows_kernel        temp chunks -> ordered CSR
//   brwt_get_kernel       K2  BRWT::get              (BRWT.cpp:9-24)
//   lane_enumerate_kernel K3a leaf call_ones         (BRWT.cpp:65-71)
//   lane_select_kernel    K3b rows[i] = select1(rows[i]+1) at one ancestor (BRWT.cpp:81-83)
#pragma once
#include "ggq_device.cuh"
#include "ggq_rows.cuh"

namespace ggq {


// ------------------------------------------------------------------------------------------
// Arena construction
// ------------------------------------------------------------------------------------------
struct ChunkTable {
    const u32 *chunk_lane;    // g*4 + lane
    const u32 *chunk_first;   // first block of the chunk
    const u32 *lane_chunk_off;  // per lane li: index of its first chunk (size 4*G + 1)
    u64 n_chunks;
};

// One CTA per chunk (kChunkBlocks threads): copy the plain bits of one block (96, or 224 in the capacity layout) into
// the slot's data words.  lane_plain_words[li] = words the lane owns in `plain`: reads beyond them are zeros.
template <bool WIDE>
__global__ void __launch_bounds__(kChunkBlocks)
pack_plain_kernel(uint4 *slots, const u64 *group_hot, const u64 *group_nblocks, ChunkTable ct,
                  const u32 *plain, const u64 *lane_plain_off, const u64 *lane_plain_words) {
    constexpr u32 DW = SlotGeom<WIDE>::kWords;
    u64 chunk = blockIdx.x + u64(blockIdx.y) * gridDim.x;
    if (chunk >= ct.n_chunks) return;
    u32 li = ct.chunk_lane[chunk];
    u32 g = li >> 2;
    u64 b = u64(ct.chunk_first[chunk]) + threadIdx.x;
    if (b >= group_nblocks[g]) return;
    u64 hot = group_hot[g];
    const u32 *src = plain + lane_plain_off[li];
    const u64 have = lane_plain_words[li];
    u32 w[DW];
    #pragma unroll
    for (u32 i = 0; i < DW; ++i) w[i] = DW * b + i < have ? src[DW * b + i] : 0u;
    SlotT<WIDE>::store_words(slots + slot_index_t<WIDE>(hot, b, li & 3), w);
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
template <bool WIDE>
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
    u32 c = 0;
    if (b < group_nblocks[g]) c = SlotT<WIDE>::load_plain(slots + slot_index_t<WIDE>(hot, b, li & 3)).popc();
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
template <bool WIDE>
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
    uint4 *p = slots + slot_index_t<WIDE>(hot, b, li & 3);
    SlotT<WIDE> s{};
    u32 c = 0;
    if (b < nb) { s = SlotT<WIDE>::load_plain(p); c = s.popc(); }
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
    s.store_header(p, u32(abs_before - sbb));
    bool internal = !(lane_info[li] & kLeafFlag) || g == 0;
    if (want_select && internal) {
        u32 *my_sel = sel + lane_sel_off[li];
        u64 first = (abs_before + (1ull << kSelectShift) - 1) >> kSelectShift;
        if ((first << kSelectShift) < abs_before + c) my_sel[first] = u32(b);
        // (a block of 96 or 224 positions holds at most one multiple of 1024)
        if (b == nb - 1) {  // sentinel sample after the last real one
            u64 ones = abs_before + c;
            my_sel[(ones + (1ull << kSelectShift) - 1) >> kSelectShift] = u32(b);
        }
    }
}

// CSR assembly in the caller's query order: 8 lanes copy one row from its chunk in the arena.  The chunk of a row sits
// where its tile happened to allocate it (processing order), so the read is a dependent random gather: every 8-lane
// group handles kGatherRows rows with all metadata loads, then all source loads, issued back to back.
constexpr int kGatherRows = 4;
template <typename OutT>
__global__ void __launch_bounds__(256)
gather_rows_kernel(const u32 *tmp_cols, const u64 *row_tmp_off, const u64 *offsets, u64 n, OutT *out) {
    const u64 grp = (u64(blockIdx.x) * blockDim.x + threadIdx.x) >> 3;
    const u32 sub = threadIdx.x & 7;
    const u64 n_grp = (u64(gridDim.x) * blockDim.x) >> 3;
    u64 dst0[kGatherRows], len[kGatherRows], src0[kGatherRows];
    #pragma unroll
    for (int k = 0; k < kGatherRows; ++k) {
        const u64 q = grp + u64(k) * n_grp;   // neighbouring groups take neighbouring rows: coalesced metadata
        dst0[k] = len[k] = src0[k] = 0;
        if (q < n) {
            dst0[k] = offsets[q];
            len[k] = offsets[q + 1] - dst0[k];
            src0[k] = row_tmp_off[q];
        }
    }
    u32 first[kGatherRows];
    #pragma unroll
    for (int k = 0; k < kGatherRows; ++k) first[k] = sub < len[k] ? tmp_cols[src0[k] + sub] : 0u;
    #pragma unroll
    for (int k = 0; k < kGatherRows; ++k) {
        if (sub < len[k]) out[dst0[k] + sub] = OutT(first[k]);
        for (u64 i = sub + 8; i < len[k]; i += 8) out[dst0[k] + i] = OutT(tmp_cols[src0[k] + i]);
    }
}

// Small batches (a scalar get_row is a batch of one): one CTA turns the descent's per-row counts and chunks into the
// final CSR and writes it STRAIGHT INTO PINNED HOST MEMORY (zero-copy stores over PCIe), so that a call costs one
// memset, two kernels and one synchronisation instead of ~14 stream operations and two round trips.
//   host[0] = total ids, host[1] = status (bit 0: descent stack overflow, bit 1: does not fit -> normal path),
//   host[2 .. 2+n] = offsets, then the u64 column ids.
constexpr int kSmallRows = 1024;
__global__ void __launch_bounds__(256)
small_finish_kernel(const u32 *row_cnt, const u64 *row_tmp_off, const u32 *tmp_cols, u64 tmp_cap, u64 *ctrl,
                    u32 n, u64 *host, u64 host_cap_ids, u64 *done) {
    __shared__ u64 s_off[kSmallRows + 1];
    __shared__ u64 s_src[kSmallRows];
    __shared__ u32 s_warp[8];
    __shared__ u64 s_carry;
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (u32 q0 = 0; q0 < n; q0 += 256) {   // block-wide exclusive scan of the counts, 256 rows at a time
        const u32 q = q0 + threadIdx.x;
        const u32 c = q < n ? row_cnt[q] : 0;
        u32 inc = c;
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const u32 v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        u32 before = 0, all = 0;
        #pragma unroll
        for (u32 w = 0; w < 8; ++w) {
            const u32 v = s_warp[w];
            if (w < warp) before += v;
            all += v;
        }
        const u64 carry = s_carry;
        if (q < n) {
            s_off[q] = carry + before + inc - c;
            s_src[q] = row_tmp_off[q];
        }
        __syncthreads();
        if (threadIdx.x == 0) s_carry = carry + all;
        __syncthreads();
    }
    const u64 total = s_carry;
    if (threadIdx.x == 0) s_off[n] = total;
    __syncthreads();
    const bool fits = total <= host_cap_ids && ctrl[0] <= tmp_cap;   // (ctrl[0]: ids the descent wanted to store)
    if (threadIdx.x == 0) {
        host[0] = total;
        host[1] = (ctrl[4] ? 1u : 0u) | (fits ? 0u : 2u);
    }
    if (fits) {
        for (u32 q = threadIdx.x; q <= n; q += 256) host[2 + q] = s_off[q];
        u64 *cols = host + 2 + n + 1;
        for (u64 p = threadIdx.x; p < total; p += 256) {
            u32 lo = 0, hi = n;   // last row with s_off[row] <= p
            while (hi - lo > 1) {
                const u32 mid = (lo + hi) >> 1;
                if (s_off[mid] <= p) lo = mid; else hi = mid;
            }
            cols[p] = tmp_cols[s_src[lo] + (p - s_off[lo])];
        }
    }
    // `done` (pinned, zeroed by the host before the launch): the control words are left cleared — the next small call on
    // this workspace needs no memset — and the flag is set once everything above is visible to the host, which may
    // poll it instead of synchronising the stream.
    if (done) {
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            #pragma unroll
            for (int i = 0; i < 9; ++i) ctrl[i] = 0;
            __threadfence_system();
            *reinterpret_cast<volatile u64 *>(done) = 1;
        }
    }
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

// get_labels(indices, ratio) / get_top_labels thresholds (annotate_static.cpp:86-91, annotate.cpp:63-66): keep the
// columns with counts[j] != 0 && counts[j] >= min_count, ascending, as (column, count) pairs.  One CTA: the
// counter array has num_columns entries (thousands), the histogram behind it had billions of increments.
__global__ void __launch_bounds__(1024)
label_filter_kernel(const u64 *counts, u32 n_cols, u64 min_count, u64 *out_cols, u64 *out_counts, u64 *n_out) {
    __shared__ u32 warp_cnt[32];
    __shared__ u32 base_s;
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) base_s = 0;
    __syncthreads();
    for (u32 j0 = 0; j0 < n_cols; j0 += 1024) {
        const u32 j = j0 + threadIdx.x;
        const u64 c = j < n_cols ? counts[j] : 0;
        const bool keep = c != 0 && c >= min_count;
        const u32 bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) warp_cnt[warp] = __popc(bal);
        __syncthreads();
        u32 before = 0, total = 0;
        for (u32 w = 0; w < 32; ++w) {
            const u32 v = warp_cnt[w];
            if (w < warp) before += v;
            total += v;
        }
        const u32 base = base_s;
        if (keep) {
            const u32 o = base + before + __popc(bal & ((1u << lane) - 1u));
            out_cols[o] = j;
            out_counts[o] = c;
        }
        __syncthreads();
        if (threadIdx.x == 0) base_s = base + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_out = base_s;
}

// Cross-query batching (the reference runs one execute_query task per sequence, main.cpp:462-504): one CTA per
// query builds the histogram of the query's column ids in shared memory (its rows are a contiguous CSR range),
// applies the get_labels threshold (annotate_static.cpp:79-91) and writes the kept (column, count) pairs,
// ascending, to a staging area at the query's CSR start (kept <= ids of the query, so the ranges never overlap).
__global__ void __launch_bounds__(256)
query_labels_kernel(const u64 *csr_offsets, const u32 *ids, const u64 *query_offsets, u32 n_cols, const u64 *min_counts,
                    u32 *stage_cols, u32 *stage_counts, u32 *kept) {
    extern __shared__ u32 s_hist[];
    __shared__ u32 warp_cnt[8];
    __shared__ u32 base_s;
    const u32 q = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const u64 r0 = query_offsets[q], r1 = query_offsets[q + 1];
    const u64 i0 = csr_offsets[r0], i1 = csr_offsets[r1];
    for (u32 j = threadIdx.x; j < n_cols; j += 256) s_hist[j] = 0;
    if (threadIdx.x == 0) base_s = 0;
    __syncthreads();
    for (u64 i = i0 + threadIdx.x; i < i1; i += 256) atomicAdd(s_hist + ids[i], 1u);
    __syncthreads();
    // min_counts[q] = ratio == 0 ? 1 : ceil(indices.size() * ratio), evaluated on the host exactly as
    // annotate_static.cpp:79-81 does
    const u64 min_count = min_counts[q];
    for (u32 j0 = 0; j0 < n_cols; j0 += 256) {
        const u32 j = j0 + threadIdx.x;
        const u32 c = j < n_cols ? s_hist[j] : 0;
        const bool keep = c != 0 && c >= min_count;
        const u32 bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) warp_cnt[warp] = __popc(bal);
        __syncthreads();
        u32 before = 0, total = 0;
        #pragma unroll
        for (u32 w = 0; w < 8; ++w) {
            const u32 v = warp_cnt[w];
            if (w < warp) before += v;
            total += v;
        }
        const u32 base = base_s;
        if (keep) {
            const u64 o = i0 + base + before + __popc(bal & ((1u << lane) - 1u));
            stage_cols[o] = j;
            stage_counts[o] = c;
        }
        __syncthreads();
        if (threadIdx.x == 0) base_s = base + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) kept[q] = base_s;
}

// Small batches of query sequences (a handful of concurrent `classify` tasks: <= 16 K indices in all): no CSR at all.
// One CTA per query builds the histogram straight from the unordered chunks the descent left in its arena
// (order is irrelevant to a histogram), thresholds it, reserves room with one atomic and writes the kept
// (column | count << 32) pairs, ascending, into PINNED HOST memory; the host waits once.
//   host[0] = status, zeroed by the host before the launch (bit 0: descent stack overflow, bit 1: arena too small)
//   host[1] = completion flag, zeroed by the host before the launch: the LAST CTA to finish (ticket in ctrl[8]) clears
//   the nine control words — the next direct call on this workspace needs no memset — and then sets it to 1, after
//   every CTA's writes have been fenced to the system: the host may poll it instead of synchronising the stream
//   host[2 + 2q], host[3 + 2q] = first pair and number of pairs of query q (first + number > capacity: the pairs
//   were not written; the host falls back to the general path);  pairs from host[2 + 2 n_queries]
// query_offsets / min_counts may live in pinned host memory (read once per CTA).
__device__ __forceinline__ void direct_kernel_epilogue(u64 *ctrl, u32 n_queries, u64 *host) {
    __threadfence_system();   // every thread: its pairs are visible to the host before the ticket is taken
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned long long t = atomicAdd(reinterpret_cast<unsigned long long *>(ctrl + 8), 1ull);
        if (t == n_queries - 1) {   // all other CTAs are past their last read of ctrl and their last write to host
            __threadfence();
            #pragma unroll
            for (int i = 0; i < 9; ++i) ctrl[i] = 0;
            __threadfence_system();
            *reinterpret_cast<volatile u64 *>(host + 1) = 1;
            __threadfence_system();
        }
    }
}

__global__ void __launch_bounds__(256)
query_labels_direct_kernel(const u32 *row_cnt, const u64 *row_tmp_off, const u32 *tmp_cols, u64 tmp_cap, u64 *ctrl,
                           const u64 *query_offsets, const u64 *min_counts, u32 n_cols, u32 n_queries, u64 *host,
                           u64 host_cap_pairs) {
    extern __shared__ u32 s_hist[];
    __shared__ u32 warp_cnt[8];
    __shared__ u32 base_s;
    __shared__ u64 out_base_s;
    const u32 q = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool overflow = *reinterpret_cast<const int *>(ctrl + 4) != 0, too_small = ctrl[0] > tmp_cap;
    if (overflow || too_small) {
        if (q == 0 && threadIdx.x == 0) { host[0] = (overflow ? 1u : 0u) | (too_small ? 2u : 0u); }
        direct_kernel_epilogue(ctrl, n_queries, host);
        return;
    }
    const u64 r0 = query_offsets[q], r1 = query_offsets[q + 1];
    for (u32 j = threadIdx.x; j < n_cols; j += 256) s_hist[j] = 0;
    __syncthreads();
    for (u64 r = r0 + threadIdx.x; r < r1; r += 256) {
        const u64 off = row_tmp_off[r];
        const u32 c = row_cnt[r];
        for (u32 i = 0; i < c; ++i) atomicAdd(s_hist + tmp_cols[off + i], 1u);
    }
    __syncthreads();
    const u64 min_count = min_counts[q];
    u32 mine = 0;
    for (u32 j = threadIdx.x; j < n_cols; j += 256) { const u32 c = s_hist[j]; mine += c != 0 && c >= min_count; }
    for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    if (lane == 0) warp_cnt[warp] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        u32 kept = 0;
        for (u32 w = 0; w < 8; ++w) kept += warp_cnt[w];
        const u64 at = atomicAdd(reinterpret_cast<unsigned long long *>(ctrl + 7), (unsigned long long)kept);
        host[2 + 2 * u64(q)] = at;
        host[3 + 2 * u64(q)] = kept;
        out_base_s = at + kept > host_cap_pairs ? ~0ull : at;
        base_s = 0;
    }
    __syncthreads();
    const u64 out_base = out_base_s;
    u64 *pairs = host + 2 + 2 * u64(n_queries);
    for (u32 j0 = 0; j0 < n_cols && out_base != ~0ull; j0 += 256) {
        const u32 j = j0 + threadIdx.x;
        const u32 c = j < n_cols ? s_hist[j] : 0;
        const bool keep = c != 0 && c >= min_count;
        const u32 bal = __ballot_sync(0xffffffffu, keep);
        __syncthreads();
        if (lane == 0) warp_cnt[warp] = __popc(bal);
        __syncthreads();
        u32 before = 0, total = 0;
        #pragma unroll
        for (u32 w = 0; w < 8; ++w) {
            const u32 v = warp_cnt[w];
            if (w < warp) before += v;
            total += v;
        }
        const u32 base = base_s;
        if (keep) pairs[out_base + base + before + __popc(bal & ((1u << lane) - 1u))] = u64(j) | (u64(c) << 32);
        __syncthreads();
        if (threadIdx.x == 0) base_s = base + total;
    }
    direct_kernel_epilogue(ctrl, n_queries, host);
}

// staging -> dense per-query lists (8 lanes copy one query)
__global__ void query_labels_gather_kernel(const u64 *csr_offsets, const u64 *query_offsets, const u64 *label_offsets,
                                           const u32 *stage_cols, const u32 *stage_counts, u64 n_queries,
                                           u64 *out_cols, u64 *out_counts) {
    const u64 q = (u64(blockIdx.x) * blockDim.x + threadIdx.x) >> 3;
    const u32 sub = threadIdx.x & 7;
    if (q >= n_queries) return;
    const u64 src = csr_offsets[query_offsets[q]];
    const u64 dst = label_offsets[q], len = label_offsets[q + 1] - dst;
    for (u64 i = sub; i < len; i += 8) {
        out_cols[dst + i] = stage_cols[src + i];
        out_counts[dst + i] = stage_counts[src + i];
    }
}

// Batch ordering of the pipelined host API: a counting sort on the top bits of the row id (<= 2^16 buckets) in small
// kernels (no shared memory to speak of, few registers).  cub's onesweep passes need 49 KB of shared memory and 30 K
// registers per CTA and cannot run beside the descent of the previous chunk; these can (the pipelined descent leaves
// one CTA slot per SM free), which takes the ordering off the critical path of ggq_get_rows_into.  The order inside
// a bucket is arrival order of the atomics: irrelevant to the result (perm maps every row back to its position),
// and as good for locality as the stable order.  The bucket counters are scanned by the block-scan kernels below
// (a single-CTA scan of 2^16 counters took 116 us; these take ~10).
__global__ void __launch_bounds__(256)
order_hist_kernel(const u64 *rows, u64 n, u32 shift, u32 last_bucket, u32 *cnt) {
    const u64 stride = u64(gridDim.x) * blockDim.x;
    for (u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        const u64 b = rows[i] >> shift;
        atomicAdd(cnt + (b < last_bucket ? u32(b) : last_bucket), 1u);
    }
}

__global__ void __launch_bounds__(256)
order_scatter_kernel(const u64 *rows, u64 n, u32 shift, u32 last_bucket, u32 *cursor, u64 *rows_sorted, u32 *perm) {
    const u64 stride = u64(gridDim.x) * blockDim.x;
    for (u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        const u64 r = rows[i];
        const u64 b = r >> shift;
        const u32 pos = atomicAdd(cursor + (b < last_bucket ? u32(b) : last_bucket), 1u);
        rows_sorted[pos] = r;
        perm[pos] = u32(i);
    }
}

// Offsets of the pipelined chunks: exclusive sum of the per-row counts (u32) to u64 offsets[0..n_rows] in three small
// kernels (block sums, scan of the block sums, block-local scan).  cub::DeviceScan's kernel does not fit beside a
// running descent either; a chunk whose scan has to wait for the NEXT chunk's descent delays the host by a whole
// descent (GGQ_TRACE timeline).
constexpr u32 kScanItems = 16, kScanBlock = 256 * kScanItems;

__device__ __forceinline__ u64 warp_inclusive_u64(u64 v, u32 lane) {
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u64 x = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += x;
    }
    return v;
}

__global__ void __launch_bounds__(256)
cnt_block_sum_kernel(const u32 *cnt, u64 n_rows, u64 *block_sum) {
    __shared__ u64 wsum[8];
    const u64 base = u64(blockIdx.x) * kScanBlock;
    u64 s = 0;
    #pragma unroll 4
    for (u32 j = 0; j < kScanItems; ++j) {
        const u64 i = base + j * 256 + threadIdx.x;
        if (i < n_rows) s += cnt[i];
    }
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    s = warp_inclusive_u64(s, lane);
    if (lane == 31) wsum[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        u64 t = 0;
        for (u32 w = 0; w < 8; ++w) t += wsum[w];
        block_sum[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(256)
block_sum_scan_kernel(u64 *block_sum, u32 n_blocks) {
    __shared__ u64 wsum[8];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u64 carry = 0;
    for (u32 b0 = 0; b0 < n_blocks; b0 += 256) {
        const u32 i = b0 + threadIdx.x;
        const u64 v = i < n_blocks ? block_sum[i] : 0;
        const u64 inc = warp_inclusive_u64(v, lane);
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        u64 pre = carry, tot = 0;
        for (u32 w = 0; w < 8; ++w) {
            if (w < warp) pre += wsum[w];
            tot += wsum[w];
        }
        if (i < n_blocks) block_sum[i] = pre + inc - v;
        carry += tot;
        __syncthreads();
    }
}

template <typename OutT>
__global__ void __launch_bounds__(256)
cnt_block_scan_kernel(const u32 *cnt, u64 n_rows, const u64 *block_off, OutT *offsets) {
    __shared__ u64 wsum[8];
    const u64 i0 = u64(blockIdx.x) * kScanBlock + u64(threadIdx.x) * kScanItems;
    u32 c[kScanItems];
    if (i0 + kScanItems <= n_rows) {
        const uint4 *p = reinterpret_cast<const uint4 *>(cnt + i0);
        #pragma unroll
        for (u32 j = 0; j < kScanItems / 4; ++j) {
            const uint4 v = p[j];
            c[4 * j] = v.x; c[4 * j + 1] = v.y; c[4 * j + 2] = v.z; c[4 * j + 3] = v.w;
        }
    } else {
        #pragma unroll
        for (u32 j = 0; j < kScanItems; ++j) c[j] = i0 + j < n_rows ? cnt[i0 + j] : 0u;
    }
    u64 sum = 0;
    #pragma unroll
    for (u32 j = 0; j < kScanItems; ++j) sum += c[j];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const u64 inc = warp_inclusive_u64(sum, lane);
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    u64 run = block_off[blockIdx.x] + inc - sum;
    for (u32 w = 0; w < warp; ++w) run += wsum[w];
    #pragma unroll
    for (u32 j = 0; j < kScanItems; ++j) {
        if (i0 + j <= n_rows) offsets[i0 + j] = OutT(run);   // n_rows + 1 offsets
        run += c[j];
    }
}

// control words -> pinned host memory by a store from the device: a cudaMemcpyAsync of 40 bytes would queue behind
// the previous chunk's 40 MB device->host copy on the copy engine and delay the host's "chunk k is done"
__global__ void publish_ctrl_kernel(const u64 *ctrl, u64 *host) {
    if (threadIdx.x < 8) host[threadIdx.x] = ctrl[threadIdx.x];
    __threadfence_system();
}

// chunk-local CSR offsets -> global offsets (pipelined host API)
__global__ void add_base_kernel(u64 *p, u64 n, u64 base) {
    u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) p[i] += base;
}

// compact host API: u32 row ids in, per-row counts narrowed to 1 / 2 / 4 bytes out (flag[0] |= 1 if one does not fit)
__global__ void widen_rows_kernel(const u32 *in, u64 n, u64 *out) {
    u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[i];
}
__global__ void offsets_to_counts_kernel(const u64 *off, u64 n, u32 *cnt) {
    u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) cnt[i] = u32(off[i + 1] - off[i]);
}
template <typename T>
__global__ void narrow_counts_kernel(const u32 *cnt, u64 n, T *out, u32 *flag) {
    u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 c = cnt[i];
    const u32 lim = sizeof(T) >= 4 ? 0xFFFFFFFFu : (1u << (8 * sizeof(T))) - 1u;
    if (c > lim) atomicOr(flag, 1u);
    out[i] = T(c < lim ? c : lim);
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

// One lane of the arena back to plain bits (3 or 7 data words per block): the inverse of pack_plain_kernel, used by
// ggq_brwt_serialize.
template <bool WIDE>
__global__ void unpack_lane_kernel(const uint4 *slots, u64 hot, u64 n_blocks, u32 lane, u32 *plain) {
    constexpr u32 DW = SlotGeom<WIDE>::kWords;
    const u64 b = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= n_blocks) return;
    const SlotT<WIDE> s = SlotT<WIDE>::load_plain(slots + slot_index_t<WIDE>(hot, b, lane));
    #pragma unroll
    for (u32 i = 0; i < DW; ++i) plain[DW * b + i] = s.word(i);
}

// ------------------------------------------------------------------------------------------
// Leaf codes (ggq_rows.cuh): for an internal node Z with first child group g0 and ones(Z) = n_pos positions in its
// children's vectors, codes[p] = DFS index (relative to the subtree's first leaf, rank0) of the only leaf set below
// Z at position p, or kCodeEscape when the subtree holds several (or, in a damaged tree, none).  One thread per
// position walks the subtree depth-first and stops at the second leaf.  `sample`: positions i * stride are examined,
// nothing is stored, *escapes counts the escape codes (the engine decides per node
// whether codes pay: DESIGN.md 1.1b).
// ------------------------------------------------------------------------------------------
constexpr int kCodeStack = 64;   // pending groups of one walk; the engine codes only subtrees with fewer groups

template <bool WIDE>
__global__ void __launch_bounds__(256)
leaf_code_kernel(Img img, u32 g0, u64 n_pos, u32 rank0, u64 stride, bool sample, unsigned char *codes, unsigned long long *escapes) {
    const u64 n_work = (n_pos + stride - 1) / stride;
    u32 esc_local = 0;
    for (u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x; i < n_work; i += u64(gridDim.x) * blockDim.x) {
        const u64 p = i * stride;
        u32 st_g[kCodeStack];
        u64 st_r[kCodeStack];
        int sp = 0;
        st_g[0] = g0; st_r[0] = p; sp = 1;
        u32 found = 0, code = kCodeEscape;
        while (sp > 0 && found < 2) {
            --sp;
            const u32 g = st_g[sp];
            const u64 r = st_r[sp];
            const u64 hot = __ldg(img.group_hot + g);
            const u32 cnt = u32(hot >> kGhCntShift) & 7u;
            if (((hot >> kGhMoreShift) & 1u) && sp < kCodeStack) { st_g[sp] = g + 1; st_r[sp] = r; ++sp; }
            u64 blk;
            u32 bit;
            SlotT<WIDE>::div(r, blk, bit);
            // children in reverse so that the walk is depth-first in child order (irrelevant to the result)
            for (int k = int(cnt) - 1; k >= 0; --k) {
                const u32 li = g * 4 + u32(k);
                const SlotT<WIDE> s = SlotT<WIDE>::load_cached(slot_addr_t<WIDE>(img, hot, blk, u32(k)));
                if (!s.test(bit)) continue;
                const u32 info = __ldg(img.lane_info + li);
                if (info & kLeafFlag) {
                    ++found;
                    code = (info & ~kLeafFlag) - rank0;
                } else if (sp < kCodeStack) {
                    st_g[sp] = info;
                    st_r[sp] = lane_rank_base_h(img, li, blk, s.header()) + s.rank_before(bit);
                    ++sp;
                } else {
                    found = 2;   // cannot happen for subtrees the engine selects; be safe: escape
                }
            }
        }
        const u32 out = (found == 1 && code < kCodeEscape) ? code : kCodeEscape;
        if (sample) esc_local += out == kCodeEscape;
        else codes[p] = (unsigned char)out;
    }
    if (sample) {
        for (int o = 16; o; o >>= 1) esc_local += __shfl_xor_sync(0xffffffffu, esc_local, o);
        if ((threadIdx.x & 31) == 0 && esc_local) atomicAdd(escapes, (unsigned long long)esc_local);
    }
}

// ------------------------------------------------------------------------------------------
// K2: get(row, column).  One thread per query walks the column's root-to-leaf lane path.
// ------------------------------------------------------------------------------------------
template <bool WIDE>
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
            u64 blk;
            u32 bit;
            if (WIDE) {
                SlotT<WIDE>::div(r, blk, bit);
            } else {
                blk = r / kSlotBits;
                bit = u32(r - blk * kSlotBits);
            }
            const SlotT<WIDE> s = SlotT<WIDE>::load(slot_addr_t<WIDE>(img, hot, blk, li & 3));
            if (!s.test(bit)) { res = false; break; }
            if (p + 1 < p1) r = lane_rank_base_h(img, li, blk, s.header()) + s.rank_before(bit);
        }
    }
    out[t] = res;
}

// ------------------------------------------------------------------------------------------
// K3: get_column
// ------------------------------------------------------------------------------------------
// a: positions of the ones of one lane, ascending (one thread per block; the rank header is the output index)
template <bool WIDE>
__global__ void lane_enumerate_kernel(Img img, u32 li, u64 *out) {
    u32 g = li >> 2;
    u64 nb = img.group_nblocks[g];
    u64 b = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= nb) return;
    u64 hot = img.group_hot[g];
    const SlotT<WIDE> s = SlotT<WIDE>::load_cached(slot_addr_t<WIDE>(img, hot, b, li & 3));
    u64 o = lane_rank_base_h(img, li, b, s.header());
    u64 pos = b * SlotGeom<WIDE>::kBits;
    #pragma unroll
    for (u32 i = 0; i < SlotGeom<WIDE>::kWords; ++i) {
        u32 w = s.word(i);
        while (w) { out[o++] = pos + 32 * i + __ffs(w) - 1; w &= w - 1; }
    }
}

// select of the q-th (0-based) one of lane li
template <bool WIDE>
__device__ __forceinline__ u64 lane_select0(const Img &img, u32 li, u64 hot, u64 q) {
    const u32 *sel = img.sel + img.lane_sel_off[li];
    u64 si = q >> kSelectShift;
    u64 lo = sel[si], hi = sel[si + 1];
    // last block in [lo, hi] whose absolute header is <= q
    while (lo < hi) {
        u64 mid = (lo + hi + 1) >> 1;
        const u32 h = __ldg(reinterpret_cast<const u32 *>(slot_addr_t<WIDE>(img, hot, mid, li & 3)));   // the rank header
        if (lane_rank_base_h(img, li, mid, h) <= q) lo = mid; else hi = mid - 1;
    }
    const SlotT<WIDE> s = SlotT<WIDE>::load_cached(slot_addr_t<WIDE>(img, hot, lo, li & 3));
    u64 before = lane_rank_base_h(img, li, lo, s.header());
    return lo * SlotGeom<WIDE>::kBits + s.select(u32(q - before));
}

// b: rows[i] = select1(rows[i] + 1) in lane li  (BRWT.cpp:81-83)
template <bool WIDE>
__global__ void lane_select_kernel(Img img, u32 li, u64 *rows, u64 n) {
    u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    u64 hot = __ldg(img.group_hot + (li >> 2));
    rows[i] = lane_select0<WIDE>(img, li, hot, rows[i]);
}

}  // namespace ggq
