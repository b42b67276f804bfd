// libggq.so — host side of the engine: image upload, workspaces, kernel launches, the C ABI.
// Kernels: ggq_kernels.cuh (queries, arena construction), ggq_builder.cuh (synthetic BRWT),
// ggq_bv.cuh (bit_vector_{stat,sd,rrr} layouts).  No CPU fallback anywhere: every compute entry
// point launches CUDA kernels or fails.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "../../include/ggq.h"
#include "ggq_format.hpp"
#include "ggq_rrr_write.hpp"
#include "ggq_tree.hpp"
#include "ggq_kernels.cuh"
#include "ggq_engine.hpp"
#include "ggq_builder.cuh"
#include "ggq_build_greedy.cuh"
#include "ggq_bv.cuh"
#include "ggq_colmajor.cuh"
#include "ggq_image.hpp"
#include "ggq_shard.cuh"

namespace ggq {

thread_local std::string g_last_error;

int fail(int code, const std::string &msg) {
    g_last_error = msg;
    return code;
}

// Slot layout of the BRWT matrices created from now on (ggq_set_default_option("wide_slots", v); GGQ_WIDE_SLOTS=1 in the
// environment sets the initial value): 0 = 16-byte slots (default), 1 = the capacity layout (ggq_layout.hpp).
static std::atomic<int> g_default_wide{-1};
u32 default_slot_bits() {
    int v = g_default_wide.load(std::memory_order_relaxed);
    if (v < 0) {
        const char *e = std::getenv("GGQ_WIDE_SLOTS");
        v = (e && *e && *e != '0') ? 1 : 0;
        g_default_wide.store(v, std::memory_order_relaxed);
    }
    return v ? kWideSlotBits : kSlotBits;
}

// ------------------------------------------------------------------------------------------
// Image upload
// ------------------------------------------------------------------------------------------
template <typename T, typename S>
static void upload_vec(DeviceBuf<T> &dst, const std::vector<S> &src, size_t min_elems = 1) {
    static_assert(sizeof(T) == sizeof(S), "element size mismatch");
    dst.alloc(std::max(src.size(), min_elems));
    if (!src.empty()) CUDA_CHECK(cudaMemcpy(dst.p, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
}

// chunk table for the header passes: one chunk = kChunkBlocks blocks of one lane
void build_chunk_table(const FlatTree &t, std::vector<u32> *chunk_lane, std::vector<u32> *chunk_first,
                       std::vector<u32> *lane_chunk_off) {
    u32 G = t.n_groups();
    lane_chunk_off->assign(size_t(G) * 4 + 1, 0);
    chunk_lane->clear();
    chunk_first->clear();
    for (u32 g = 0; g < G; ++g) {
        u32 cnt = u32(t.group_hot[g] >> kGhCntShift) & 7u;
        u64 nb = t.group_nblocks[g];
        if (nb > 0xFFFFFFFFull) throw ParseError("vector has more blocks than the engine supports");
        for (u32 k = 0; k < 4; ++k) {
            (*lane_chunk_off)[g * 4 + k] = u32(chunk_lane->size());
            if (k >= cnt) continue;
            for (u64 b = 0; b < nb; b += kChunkBlocks) {
                chunk_lane->push_back(g * 4 + k);
                chunk_first->push_back(u32(b));
            }
        }
    }
    (*lane_chunk_off)[size_t(G) * 4] = u32(chunk_lane->size());
    if (chunk_lane->size() > 0xFFFFFFF0ull) throw ParseError("too many chunks");
}

static dim3 chunk_grid(u64 n_chunks) {
    u64 x = std::min<u64>(n_chunks, 1u << 20);
    u64 y = (n_chunks + x - 1) / std::max<u64>(x, 1);
    return dim3(unsigned(std::max<u64>(x, 1)), unsigned(std::max<u64>(y, 1)));
}

// Builds rank headers, superblock bases, lane totals and select samples over a filled arena.
void Matrix::finalize_arena(const FlatTree &t, bool ones_known) {
    std::vector<u32> chunk_lane, chunk_first, lane_chunk_off;
    build_chunk_table(t, &chunk_lane, &chunk_first, &lane_chunk_off);
    DeviceBuf<u32> d_cl, d_cf, d_lco, d_sum;
    DeviceBuf<u64> d_base;
    upload_vec(d_cl, chunk_lane);
    upload_vec(d_cf, chunk_first);
    upload_vec(d_lco, lane_chunk_off);
    u64 nch = chunk_lane.size();
    d_sum.alloc(std::max<u64>(nch, 1));
    d_base.alloc(std::max<u64>(nch, 1));
    ChunkTable ct{d_cl.p, d_cf.p, d_lco.p, nch};
    u32 n_lanes = t.n_groups() * 4;
    if (nch) {
        if (wide()) chunk_sum_kernel<true><<<chunk_grid(nch), kChunkBlocks>>>(slots.p, group_hot.p, group_nblocks.p, ct, d_sum.p);
        else chunk_sum_kernel<false><<<chunk_grid(nch), kChunkBlocks>>>(slots.p, group_hot.p, group_nblocks.p, ct, d_sum.p);
        lane_scan_kernel<<<n_lanes, 256>>>(d_lco.p, d_sum.p, d_base.p, lane_ones.p, lane_sb_off.p, sb_base.p, n_lanes);
        CUDA_CHECK(cudaGetLastError());
    }
    if (!ones_known) {
        // synthetic trees: select sample geometry depends on the ones just counted
        std::vector<u64> ones(n_lanes);
        CUDA_CHECK(cudaMemcpy(ones.data(), lane_ones.p, n_lanes * 8, cudaMemcpyDeviceToHost));
        FlatTree &mt = const_cast<FlatTree &>(t);
        u64 sel_total = 0;
        for (u32 li = 0; li < n_lanes; ++li) {
            u32 node = mt.lane_node[li];
            mt.lane_sel_off[li] = sel_total;
            if (node == ~0u || !mt.present(node)) continue;
            mt.nodes[node].ones = ones[li];
            if (!mt.nodes[node].children.empty() || mt.nodes[node].parent < 0)
                sel_total += ((ones[li] + (1ull << kSelectShift) - 1) >> kSelectShift) + 1;
        }
        mt.total_sel = sel_total;
        upload_vec(lane_sel_off, mt.lane_sel_off);
        sel.alloc(std::max<u64>(sel_total, 1));
    }
    if (nch) {
        if (wide()) header_kernel<true><<<chunk_grid(nch), kChunkBlocks>>>(slots.p, group_hot.p, group_nblocks.p, ct, d_base.p,
                                                                          lane_info.p, lane_sel_off.p, sel.p, 1);
        else header_kernel<false><<<chunk_grid(nch), kChunkBlocks>>>(slots.p, group_hot.p, group_nblocks.p, ct, d_base.p,
                                                                    lane_info.p, lane_sel_off.p, sel.p, 1);
        CUDA_CHECK(cudaGetLastError());
    }
    CUDA_CHECK(cudaDeviceSynchronize());
}

// Group records of the descent kernel (ggq_rows.cuh: load_rec), with the leaf-code marks of the coded lanes when
// leaf codes are in use.
void Matrix::upload_records(const FlatTree &t) {
    const bool codes_on = use_codes && !t.lane_coded.empty();
    auto is_coded = [&](u32 li) { return codes_on && t.lane_coded[li] != 0; };
    {
        std::vector<uint4> rec(size_t(t.n_groups()) * 2);
        for (u32 g = 0; g < t.n_groups(); ++g)
            for (u32 h = 0; h < 2; ++h) {
                u32 c[2];
                for (u32 k = 0; k < 2; ++k) {
                    const u32 li = 4 * g + 2 * h + k;
                    c[k] = t.lane_info[li] | (is_coded(li) ? kCodedFlag : 0u);   // (coded lanes are internal: bit 31 clear)
                }
                rec[2 * g + h] = make_uint4(u32(t.group_hot[g]), u32(t.group_hot[g] >> 32), c[0], c[1]);
            }
        upload_vec(group_rec, rec);
    }
    {   // compact 16-byte records: 21-bit children (leaf bit 20), base in units of 4 slots, coded mask in y[28..31]
        compact_rec = t.leaf_col.size() < (1u << 20) - 1 && t.n_groups() < (1u << 20) && !std::getenv("GGQ_WIDE_REC");
        std::vector<uint4> rec16(compact_rec ? size_t(t.n_groups()) : 0);
        for (u32 g = 0; compact_rec && g < t.n_groups(); ++g) {
            const u64 hot = t.group_hot[g];
            const u64 base = hot & kGhBaseMask;
            if ((base & 3) || (base >> 2) > 0xFFFFFFFFull) { compact_rec = false; break; }
            u32 c[4] = {0, 0, 0, 0};
            u32 cmask = 0;
            for (u32 k = 0; k < 4; ++k) {
                const u32 info = t.lane_info[4 * g + k];
                if (info == kRemoteLane) c[k] = (1u << 21) - 1u;
                else if (info & kLeafFlag) c[k] = (1u << 20) | (info & ~kLeafFlag);
                else c[k] = info;
                if (info != kRemoteLane && (info & ~kLeafFlag) >= (1u << 20) - 1) { compact_rec = false; break; }
                if (is_coded(4 * g + k)) cmask |= 1u << k;
            }
            const u32 flags = u32(hot >> kGhStrideShift) & 0x7fu;
            rec16[g] = make_uint4(u32(base >> 2), flags | (c[0] << 7) | (cmask << 28), c[1] | (c[2] << 21), (c[2] >> 11) | (c[3] << 10));
        }
        if (compact_rec) upload_vec(group_rec16, rec16);
        else group_rec16.release();
    }
}

void Matrix::upload_tables(const FlatTree &t) {
    upload_vec(group_hot, t.group_hot);
    upload_vec(group_nblocks, t.group_nblocks);
    upload_records(t);
    upload_vec(lane_info, t.lane_info);
    upload_vec(lane_aux, t.lane_aux);
    upload_vec(lane_sb_off, t.lane_sb_off);
    upload_vec(lane_sel_off, t.lane_sel_off);
    upload_vec(leaf_col, t.leaf_col);
    upload_vec(col_path_off, t.col_path_off);
    upload_vec(col_path, t.col_path);
    lane_ones.alloc(size_t(t.n_groups()) * 4);
    CUDA_CHECK(cudaMemset(lane_ones.p, 0, size_t(t.n_groups()) * 4 * 8));
    sb_base.alloc(std::max<u64>(t.total_sb, 1));
    CUDA_CHECK(cudaMemset(sb_base.p, 0, std::max<u64>(t.total_sb, 1) * 8));
}

void Matrix::refresh_img() {
    img.slots = slots.p;
    img.group_hot = group_hot.p;
    img.group_rec = group_rec.p;
    img.group_rec16 = compact_rec ? group_rec16.p : nullptr;
    img.group_nblocks = group_nblocks.p;
    img.lane_info = lane_info.p;
    img.lane_aux = lane_aux.p;
    img.lane_sb_off = lane_sb_off.p;
    img.lane_sel_off = lane_sel_off.p;
    img.lane_ones = lane_ones.p;
    img.sb_base = sb_base.p;
    img.sel = sel.p;
    img.leaf_col = leaf_col.p;
    img.col_path_off = col_path_off.p;
    img.col_path = col_path.p;
    img.codes = codes.p;
    img.code_rec = code_rec.p;
    img.n_rows = tree.n_rows;
    img.n_cols = u32(tree.n_cols);
    img.n_groups = tree.n_groups();
    img.multi_leaf = tree.multi_column_leaves;
}

// The instantiation of K1 for a tree: stack spill to global memory, leaves holding several columns.
template <bool WIDE>
static const void *rows_kernel_fn_t(bool spill, bool multi_leaf, bool compact) {
    if (compact) {
        if (spill) return multi_leaf ? (const void *)brwt_rows_kernel<true, true, true, WIDE> : (const void *)brwt_rows_kernel<true, false, true, WIDE>;
        return multi_leaf ? (const void *)brwt_rows_kernel<false, true, true, WIDE> : (const void *)brwt_rows_kernel<false, false, true, WIDE>;
    }
    if (spill) return multi_leaf ? (const void *)brwt_rows_kernel<true, true, false, WIDE> : (const void *)brwt_rows_kernel<true, false, false, WIDE>;
    return multi_leaf ? (const void *)brwt_rows_kernel<false, true, false, WIDE> : (const void *)brwt_rows_kernel<false, false, false, WIDE>;
}
static const void *rows_kernel_fn(bool spill, bool multi_leaf, bool compact, bool wide) {
    return wide ? rows_kernel_fn_t<true>(spill, multi_leaf, compact) : rows_kernel_fn_t<false>(spill, multi_leaf, compact);
}

// Descent kernel configuration: rows per warp batch, stack capacity, grid, query ordering.
void Matrix::configure_rows(u32 rb_request, u32 sort_request) {
    // descent kernel configuration
    bm_words = u32((tree.leaf_col.size() + 31) / 32);
    if (bm_words == 0) bm_words = 1;
    int max_smem = 0;
    CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    CUDA_CHECK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device));
    // rows per warp batch: as many as fit a ~5.5 KB/warp bitmap budget, at most 16 (kItemTBits)
    auto env_u32 = [](const char *name, u32 dflt) {
        const char *v = std::getenv(name);
        return v && *v ? u32(std::strtoul(v, nullptr, 10)) : dflt;
    };
    rb = std::min<u32>(std::max<u32>(rb_request ? rb_request : env_u32("GGQ_RB", 8), 1), kMaxRowsPerWarp);
    while (rb > 1 && size_t(rb) * bm_words * 4 > 5632) rb >>= 1;
    for (;;) {
        u32 worst = simulate_stack(tree, rb, kItemsPerIter) + 8;
        cap_s = (std::max<u32>(worst, 32) + 1) & ~1u;
        cap_spill = 0;
        // deep/wide trees: overflow to global memory (GGQ_STACK_CAP lowers the shared part: the tests use it to run
        // the SPILL instantiation on ordinary trees)
        const u32 cap_limit = std::min<u32>(std::max<u32>(env_u32("GGQ_STACK_CAP", 512) & ~1u, 16), 512);
        if (cap_s > cap_limit) { cap_spill = cap_s - cap_limit; cap_s = cap_limit; }
        rows_smem = rows_smem_bytes(cap_s, rb, bm_words);
        if (rows_smem <= size_t(max_smem) || rb == 1) break;
        rb >>= 1;
    }
    if (rows_smem > size_t(max_smem)) throw LimitError("matrix has too many columns for the descent kernel's shared-memory bitmaps");
    int per_sm = 0;
    const void *fn = rows_kernel_fn(cap_spill != 0, tree.multi_column_leaves != 0, compact_rec, wide());
    CUDA_CHECK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));  // per function, shared by all matrices
    CUDA_CHECK(cudaFuncSetAttribute(rows_kernel_fn(true, tree.multi_column_leaves != 0, compact_rec, wide()),
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));   // the safe fallback below
    {   // provable stack bound: the items of one row that are pending at any time head disjoint subtrees of the group
        // tree (children groups + continuation group), and every such subtree contains a group without successors
        u64 leaves = 0;
        const u32 G = tree.n_groups();
        for (u32 g = 0; g < G; ++g) {
            const u64 h = tree.group_hot[g];
            bool succ = ((h >> kGhMoreShift) & 1u) != 0;
            const u32 cnt = u32(h >> kGhCntShift) & 7u;
            for (u32 k = 0; k < cnt && !succ; ++k) {
                const u32 info = tree.lane_info[g * 4 + k];
                succ = !(info & kLeafFlag);   // (kRemoteLane has the leaf bit: never descended)
            }
            leaves += !succ;
        }
        cap_safe = u32(std::min<u64>(std::max<u64>(leaves, 1) * rb + 2 * kItemsPerIter, 0x7fffffffu));
        if (env_u32("GGQ_STACK_UNDERESTIMATE", 0) && cap_spill == 0)   // tests: force the overflow -> safe-mode path
            cap_s = std::max<u32>(rb + 2, std::min<u32>(cap_s, env_u32("GGQ_STACK_UNDERESTIMATE", 0)) & ~1u);
    }
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kRowsWarps * 32, rows_smem));
    if (cap_spill == 0 && !env_u32("GGQ_STACK_UNDERESTIMATE", 0)) {
        // The 8 entries of slack above the replayed stack height are a courtesy, not a guarantee (the kernel checks):
        // give some of them up when that is what keeps one more CTA resident per SM (C4: 6 instead of 5).
        const u32 worst = (std::max<u32>(simulate_stack(tree, rb, kItemsPerIter) + 2, 32) + 1) & ~1u;
        int occ_tight = per_sm;
        if (worst < cap_s)
            CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_tight, fn, kRowsWarps * 32, rows_smem_bytes(worst, rb, bm_words)));
        while (occ_tight > per_sm && cap_s > worst) {
            cap_s -= 2;
            rows_smem = rows_smem_bytes(cap_s, rb, bm_words);
            CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kRowsWarps * 32, rows_smem));
        }
    }
    if (per_sm < 1) per_sm = 1;
    rows_grid = u32(sm_count * per_sm);
    rows_grid_p = rows_grid;
    rows_smem_p = rows_smem;
    if (per_sm >= 3 && !env_u32("GGQ_NO_PIPE_ROOM", 0)) {
        size_t s = rows_smem;
        int occ = per_sm;
        while (occ >= per_sm && s + 512 <= size_t(max_smem)) {
            s += 512;
            CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, kRowsWarps * 32, s));
        }
        if (occ == per_sm - 1) { rows_smem_p = s; rows_grid_p = u32(sm_count * occ); }
    }
    // root fast path + continuation groups (see RowsArgs)
    root_child = kNoRootFast;
    root_all_ones = 0;
    any_more = 0;
    for (u64 h : tree.group_hot) any_more |= u32((h >> kGhMoreShift) & 1u);
    if (!tree.group_hot.empty()) {
        const u64 h0 = tree.group_hot[0];
        const u32 info0 = tree.lane_info[0];
        if (((h0 >> kGhCntShift) & 7u) == 1 && !((h0 >> kGhMoreShift) & 1u) && !(info0 & kLeafFlag) &&
            !env_u32("GGQ_NO_ROOT_FAST", 0)) {
            root_child = info0;
            root_sl = u32(h0 >> kGhStrideShift) & 3u;
            root_multi_sb = u32((h0 >> kGhSuperShift) & 1u);
            root_base = h0 & kGhBaseMask;
            root_all_ones = !lane_ones_h.empty() && !lane_size_h.empty() && lane_size_h[0] > 0 && lane_ones_h[0] == lane_size_h[0] &&
                            !env_u32("GGQ_NO_ROOT_ALL_ONES", 0);
        }
    }
    // query ordering (see K1): sort on the top `sort_bits` bits of the row id
    row_bits = 1;
    while (row_bits < 64 && (u64(1) << row_bits) < tree.n_rows) ++row_bits;
    sort_bits = std::min<u32>(sort_request != ~0u ? sort_request : env_u32("GGQ_SORT_BITS", 16), row_bits);
    sort_min_n = env_u32("GGQ_SORT_MIN", 4096);
}

void Matrix::derive_stats() {
    if (!tree.nodes.empty()) {   // (a matrix restored from an image file has no host nodes: stats come from the file)
        n_nodes = tree.nodes.size();
        total_column_size = 0;
        num_relations = 0;
        for (const FlatNode &n : tree.nodes) {
            total_column_size += n.size;
            if (n.children.empty())   // column-sharded: owned columns only
                for (u32 c : n.cols)
                    if (tree.col_owned.empty() || tree.col_owned[c]) num_relations += n.ones;
        }
        // per lane: length and popcount of its vector (get_column: leaf size, all-ones ancestors BRWT.cpp:77-78)
        lane_size_h.assign(tree.lane_node.size(), 0);
        lane_ones_h.assign(tree.lane_node.size(), 0);
        for (size_t li = 0; li < tree.lane_node.size(); ++li) {
            const u32 node = tree.lane_node[li];
            if (node == ~0u || !tree.present(node)) continue;
            lane_size_h[li] = tree.nodes[node].size;
            lane_ones_h[li] = tree.nodes[node].ones;
        }
    }
    configure_rows(0, ~0u);
    device_bytes = slots.bytes() + codes.bytes() * (code_bytes != 0) + code_rec.bytes() * (code_bytes != 0) +
                   sel.bytes() + sb_base.bytes() + group_hot.bytes() + group_nblocks.bytes() +
                   lane_info.bytes() * 3 + lane_sel_off.bytes() * 2 + leaf_col.bytes() + col_path.bytes() + col_path_off.bytes();
}

// Leaf codes (ggq_rows.cuh).  Which nodes get codes is decided here, top-down: a node Z qualifies
// when its subtree has at most 255 single-column leaves and fewer than kCodeStack groups; its escape rate e(Z)
// (positions with several labels below Z) is sampled on the device, and codes pay when they save probes:
//   gain(Z) = C(Z) - (1 + e(Z) * C(Z)),  C(Z) = expected probes of the plain descent below Z per visit
//           = sum over internal Y in subtree(Z) of ones(Y) * groups(Y) / ones(Z)      (uniformly random rows).
// A node that does not pay hands the question to its children.  Accepted nodes are kept in order of gain while their
// code bytes (ones(Z) each) fit the budget: GGQ_CODE_BUDGET_MB, default half the slot arena, never more than the
// free device memory minus a reserve for workspaces.
void Matrix::build_leaf_codes() {
    tree.lane_coded.clear();
    tree.code_rec.clear();
    codes.release();
    code_rec.release();
    code_bytes = 0; n_coded = 0; code_escape_rate = 0;
    auto &N = tree.nodes;
    const char *off = std::getenv("GGQ_NO_CODES");
    if (kind != BRWT || N.size() < 2 || tree.multi_column_leaves || tree.n_shards > 1 || (off && *off && *off != '0')) {
        refresh_img();
        return;
    }
    CUDA_CHECK(cudaSetDevice(device));
    const size_t nn = N.size();
    std::vector<u32> sub_groups(nn, 0), first_rank(nn, 0);
    std::vector<double> probe_mass(nn, 0.0);   // sum over internal Y in the subtree of ones(Y) * groups(Y)
    for (size_t id = nn; id-- > 0;) {          // BFS ids: children have larger ids than their parent
        const FlatNode &n = N[id];
        if (n.children.empty()) { first_rank[id] = n.leaf_rank; continue; }
        sub_groups[id] = n.n_child_groups;
        probe_mass[id] = double(n.ones) * n.n_child_groups;
        first_rank[id] = first_rank[n.children[0]];
        for (int c : n.children) { sub_groups[id] += sub_groups[c]; probe_mass[id] += probe_mass[c]; }
    }
    const double min_gain = 0.5;
    struct Pick { u32 id; double gain, esc; };
    std::vector<Pick> picks;
    DeviceBuf<unsigned long long> d_esc;
    d_esc.alloc(1);
    std::vector<int> todo;
    for (size_t i = N[0].children.size(); i-- > 0;) todo.push_back(N[0].children[i]);   // (the root itself is never coded)
    while (!todo.empty()) {
        const int id = todo.back();
        todo.pop_back();
        const FlatNode &n = N[id];
        if (n.children.empty() || !tree.present(u32(id))) continue;
        bool taken = false;
        if (n.cols.size() <= kCodeEscape && sub_groups[id] < u32(kCodeStack) && n.ones > 0) {
            const double C = probe_mass[id] / double(n.ones);
            if (C - 1.0 >= min_gain) {   // (even with no escapes at all a 1-probe subtree gains nothing)
                const u64 samples = std::min<u64>(n.ones, u64(1) << 16);
                const u64 stride = std::max<u64>(1, n.ones / samples);
                const u64 n_work = (n.ones + stride - 1) / stride;
                CUDA_CHECK(cudaMemset(d_esc.p, 0, 8));
                if (wide()) leaf_code_kernel<true><<<u32((n_work + 255) / 256), 256>>>(img, n.first_child_group, n.ones, first_rank[id], stride, true, nullptr, d_esc.p);
                else leaf_code_kernel<false><<<u32((n_work + 255) / 256), 256>>>(img, n.first_child_group, n.ones, first_rank[id], stride, true, nullptr, d_esc.p);
                CUDA_CHECK(cudaGetLastError());
                unsigned long long esc = 0;
                CUDA_CHECK(cudaMemcpy(&esc, d_esc.p, 8, cudaMemcpyDeviceToHost));
                const double e = double(esc) / double(n_work);
                const double gain = C - (1.0 + e * C);
                if (gain >= min_gain) { picks.push_back({u32(id), gain, e}); taken = true; }
            }
        }
        if (!taken)
            for (size_t i = n.children.size(); i-- > 0;) todo.push_back(n.children[i]);
    }
    if (picks.empty()) { refresh_img(); return; }
    // budget
    size_t free_b = 0, total_b = 0;
    CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
    u64 budget = slots.bytes() / 2;
    if (const char *b = std::getenv("GGQ_CODE_BUDGET_MB")) budget = std::strtoull(b, nullptr, 10) << 20;
    const u64 reserve = u64(12) << 30;
    budget = std::min<u64>(budget, free_b > reserve ? free_b - reserve : 0);
    std::stable_sort(picks.begin(), picks.end(), [](const Pick &a, const Pick &b) { return a.gain > b.gain; });
    std::vector<Pick> kept;
    u64 bytes = 0;
    for (const Pick &p : picks) {
        const u64 need = (N[p.id].ones + 63) & ~u64(63);
        if (bytes + need > budget) continue;
        bytes += need;
        kept.push_back(p);
    }
    if (kept.empty()) { refresh_img(); return; }
    codes.alloc(bytes);
    tree.lane_coded.assign(size_t(tree.n_groups()) * 4, 0);
    tree.code_rec.assign(size_t(tree.n_groups()) * 16, 0);
    int sms = 148;
    CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    u64 at = 0;
    double visits = 0, esc_visits = 0;
    for (const Pick &p : kept) {
        const FlatNode &n = N[p.id];
        const u32 li = n.group * 4 + n.lane;
        const u32 grid = u32(std::min<u64>((n.ones + 255) / 256, u64(sms) * 32));
        if (wide()) leaf_code_kernel<true><<<grid, 256>>>(img, n.first_child_group, n.ones, first_rank[p.id], 1, false, codes.p + at, nullptr);
        else leaf_code_kernel<false><<<grid, 256>>>(img, n.first_child_group, n.ones, first_rank[p.id], 1, false, codes.p + at, nullptr);
        CUDA_CHECK(cudaGetLastError());
        tree.lane_coded[li] = 1;
        tree.code_rec[4 * size_t(li) + 0] = u32(at);
        tree.code_rec[4 * size_t(li) + 1] = u32(at >> 32);
        tree.code_rec[4 * size_t(li) + 2] = first_rank[p.id];
        tree.code_rec[4 * size_t(li) + 3] = u32(n.cols.size());
        at += (n.ones + 63) & ~u64(63);
        visits += double(n.ones);
        esc_visits += double(n.ones) * p.esc;
    }
    CUDA_CHECK(cudaDeviceSynchronize());
    code_bytes = bytes;
    n_coded = u32(kept.size());
    code_escape_rate = visits > 0 ? esc_visits / visits : 0;
    upload_codes_tables();
}

// code_rec + group records after tree.lane_coded / tree.code_rec changed (or use_codes was switched)
void Matrix::upload_codes_tables() {
    if (!tree.code_rec.empty()) {
        code_rec.alloc(tree.code_rec.size() / 4);
        CUDA_CHECK(cudaMemcpy(code_rec.p, tree.code_rec.data(), tree.code_rec.size() * 4, cudaMemcpyHostToDevice));
    }
    upload_records(tree);
    refresh_img();
}

// Fills the concatenated plain bits of all present lanes on the device: d_plain (u32 words, zeroed),
// lane_plain_off[li] = first word of lane li, which owns (size + 31) / 32 + 4 words.
typedef std::function<void(const FlatTree &, u32 *, const std::vector<u64> &)> PlainFill;

// tree with plain bits per node (host-parsed: FlatNode::bits; device-built: `fill`) -> device image
static Matrix *upload_brwt(FlatTree &&tree, int device, u32 shard = 0, u32 n_shards = 1, const PlainFill *fill = nullptr) {
    std::unique_ptr<Matrix> m(new Matrix());
    m->kind = Matrix::BRWT;
    m->device = device;
    CUDA_CHECK(cudaSetDevice(device));
    m->tree = std::move(tree);
    FlatTree &t = m->tree;
    t.slot_bits = default_slot_bits();
    assign_groups(&t);
    plan_shards(&t, shard, n_shards);
    assign_geometry(&t);
    m->upload_tables(t);
    m->slots.alloc(std::max<u64>(t.total_slots, 4));
    CUDA_CHECK(cudaMemset(m->slots.p, 0, m->slots.bytes()));
    m->sel.alloc(std::max<u64>(t.total_sel, 1));
    // plain bits of all nodes, concatenated
    u32 n_lanes = t.n_groups() * 4;
    std::vector<u64> lane_plain_off(n_lanes, 0), lane_plain_words(n_lanes, 0);
    u64 total_words = 0;
    for (u32 li = 0; li < n_lanes; ++li) {
        if (t.lane_node[li] == ~0u || !t.present(t.lane_node[li])) continue;
        lane_plain_off[li] = total_words;
        const FlatNode &nd = t.nodes[t.lane_node[li]];
        lane_plain_words[li] = (fill || nd.lazy) ? (nd.size + 31) / 32 + 4 : nd.bits.w.size();
        total_words += lane_plain_words[li];
    }
    {
        DeviceBuf<u32> d_plain;
        DeviceBuf<u64> d_off, d_have;
        d_plain.alloc(std::max<u64>(total_words, 1));
        if (fill) {
            CUDA_CHECK(cudaMemset(d_plain.p, 0, d_plain.bytes()));
            (*fill)(t, d_plain.p, lane_plain_off);
        } else {
            // Lazily parsed vectors are decoded in batches of at most 2^33 plain bits (1 GiB; a longer vector is a batch
            // of its own) on the host's threads, uploaded, and forgotten: the host never holds more plain bits than that.
            const u64 kBatchBits = u64(1) << 33;
            std::vector<u32> lanes;
            for (u32 li = 0; li < n_lanes; ++li)
                if (t.lane_node[li] != ~0u && t.present(t.lane_node[li])) lanes.push_back(li);
            for (size_t i = 0; i < lanes.size();) {
                size_t j = i;
                u64 bits = 0;
                std::vector<FlatNode *> batch;
                for (; j < lanes.size(); ++j) {
                    FlatNode &nd = t.nodes[t.lane_node[lanes[j]]];
                    if (!nd.lazy) continue;
                    if (!batch.empty() && bits + nd.size > kBatchBits) break;
                    batch.push_back(&nd);
                    bits += nd.size;
                }
                decode_nodes_parallel(batch);
                for (; i < j; ++i) {
                    const u32 li = lanes[i];
                    FlatNode &nd = t.nodes[t.lane_node[li]];
                    if (nd.lazy) {
                        if (nd.bits.size != nd.size || (!nd.children.empty() && nd.bits.ones != nd.ones))
                            throw ParseError("BRWT: child length differs from the parent's popcount");
                        nd.ones = nd.bits.ones;
                        nd.raw.clear();
                    }
                    const auto &w = nd.bits.w;
                    CUDA_CHECK(cudaMemcpy(d_plain.p + lane_plain_off[li], w.data(), w.size() * 4, cudaMemcpyHostToDevice));
                    if (nd.lazy) { nd.bits.w.clear(); nd.bits.w.shrink_to_fit(); }
                }
            }
        }
        upload_vec(d_off, lane_plain_off);
        upload_vec(d_have, lane_plain_words);
        std::vector<u32> chunk_lane, chunk_first, lane_chunk_off;
        build_chunk_table(t, &chunk_lane, &chunk_first, &lane_chunk_off);
        DeviceBuf<u32> d_cl, d_cf, d_lco;
        upload_vec(d_cl, chunk_lane);
        upload_vec(d_cf, chunk_first);
        upload_vec(d_lco, lane_chunk_off);
        ChunkTable ct{d_cl.p, d_cf.p, d_lco.p, chunk_lane.size()};
        if (ct.n_chunks) {
            if (m->wide())
                pack_plain_kernel<true><<<chunk_grid(ct.n_chunks), kChunkBlocks>>>(m->slots.p, m->group_hot.p, m->group_nblocks.p,
                                                                                   ct, d_plain.p, d_off.p, d_have.p);
            else
                pack_plain_kernel<false><<<chunk_grid(ct.n_chunks), kChunkBlocks>>>(m->slots.p, m->group_hot.p, m->group_nblocks.p,
                                                                                    ct, d_plain.p, d_off.p, d_have.p);
            CUDA_CHECK(cudaGetLastError());
        }
        CUDA_CHECK(cudaDeviceSynchronize());
    }
    for (FlatNode &n : t.nodes) { n.bits.w.clear(); n.bits.w.shrink_to_fit(); n.raw.clear(); }
    m->finalize_arena(t, true);
    m->refresh_img();
    m->build_leaf_codes();
    m->derive_stats();
    return m.release();
}

// Builder tree -> FlatTree::nodes in level order (children in stored order); bn_of[flat id] = builder node.
static void shape_to_flat(const std::vector<BuildNode> &N, int root, u64 n_rows, u32 n_cols, FlatTree *t, std::vector<int> *bn_of) {
    struct Pending { int bn, parent, child_index; };
    std::deque<Pending> q;
    q.push_back({root, -1, 0});
    t->nodes.clear();
    bn_of->clear();
    while (!q.empty()) {
        const Pending cu = q.front();
        q.pop_front();
        const int id = int(t->nodes.size());
        t->nodes.emplace_back();
        bn_of->push_back(cu.bn);
        FlatNode &fn = t->nodes.back();
        fn.parent = cu.parent;
        fn.child_index = cu.child_index;
        fn.depth = cu.parent < 0 ? 0 : t->nodes[cu.parent].depth + 1;
        fn.cols = N[cu.bn].cols;
        if (cu.parent < 0)   // the root's local columns ARE the global ones (as flatten_parsed has it)
            for (u32 c = 0; c < n_cols; ++c) fn.cols[c] = c;
        fn.ones = N[cu.bn].ones;
        fn.size = cu.parent < 0 ? n_rows : t->nodes[cu.parent].ones;
        if (cu.parent >= 0) t->nodes[cu.parent].children.push_back(id);
        for (size_t i = 0; i < N[cu.bn].children.size(); ++i) q.push_back({N[cu.bn].children[i], id, int(i)});
    }
    t->n_rows = n_rows;
    t->n_cols = n_cols;
}

// BRWTBottomUpBuilder::build + BRWTOptimizer::relax on the device (ggq_build_greedy.cuh) from explicit columns:
// col_words = n_cols columns of ceil(n_rows / 64) LSB-first u64 words each.  partitioner: 0 = basic(arity), 1 = greedy.
static Matrix *build_brwt_columns(const u64 *col_words, u64 n_rows, u32 n_cols, int partitioner, u32 arity,
                                  u64 relax_arity, int device, std::vector<BuildNode> *shape_out = nullptr, int *root_out = nullptr) {
    CUDA_CHECK(cudaSetDevice(device));
    FlatTree t;
    if (n_cols == 0) {   // BRWT(): a single empty node
        t.nodes.emplace_back();
        t.n_rows = 0; t.n_cols = 0;
        return upload_brwt(std::move(t), device);
    }
    const u64 W = (n_rows + 63) / 64;
    std::vector<BuildNode> N(n_cols);
    std::vector<DeviceBuf<u64>> full(n_cols), sub_own;
    std::vector<const u64 *> sub(n_cols, nullptr);
    for (u32 j = 0; j < n_cols; ++j) {
        N[j].cols = {j};
        full[j].alloc(std::max<u64>(W, 1));
        if (W) {
            CUDA_CHECK(cudaMemcpy(full[j].p, col_words + size_t(j) * W, W * 8, cudaMemcpyHostToDevice));
            if (n_rows & 63) {   // clear padding
                const u64 last = col_words[size_t(j) * W + W - 1] & ((1ull << (n_rows & 63)) - 1ull);
                CUDA_CHECK(cudaMemcpy(full[j].p + W - 1, &last, 8, cudaMemcpyHostToDevice));
            }
        }
    }
    // sampled rows for the similarity estimates (all rows when n_rows <= 10^6)
    u64 Ws = W;
    DeviceBuf<u64> d_idx;
    bool sampled = false;
    if (partitioner == 1 && n_rows > kBuildRowsSampled) {
        const std::vector<u64> idx = build_row_sample(n_rows);
        upload_vec(d_idx, idx);
        Ws = (idx.size() + 63) / 64;
        sampled = true;
        sub_own.resize(n_cols);
        for (u32 j = 0; j < n_cols; ++j) {
            sub_own[j].alloc(Ws);
            sample_rows_kernel<<<u32((Ws + 255) / 256), 256>>>(full[j].p, d_idx.p, idx.size(), sub_own[j].p);
            sub[j] = sub_own[j].p;
        }
        CUDA_CHECK(cudaGetLastError());
    } else {
        for (u32 j = 0; j < n_cols; ++j) sub[j] = full[j].p;
    }
    std::vector<int> cur(n_cols);
    for (u32 j = 0; j < n_cols; ++j) cur[j] = int(j);
    DeviceBuf<u32> d_G;
    DeviceBuf<const u64 *> d_ptrs;
    while (cur.size() > 1) {
        const u32 k_n = u32(cur.size());
        std::vector<std::vector<u32>> part;
        if (partitioner == 1) {
            std::vector<const u64 *> ptrs(k_n);
            for (u32 i = 0; i < k_n; ++i) ptrs[i] = sub[cur[i]];
            d_ptrs.alloc(k_n);
            CUDA_CHECK(cudaMemcpy(d_ptrs.p, ptrs.data(), size_t(k_n) * sizeof(const u64 *), cudaMemcpyHostToDevice));
            d_G.alloc(size_t(k_n) * k_n);
            const u32 tiles = (k_n + kGramTile - 1) / kGramTile;
            gram_kernel<<<dim3(tiles, tiles), 256>>>(d_ptrs.p, k_n, Ws, d_G.p);
            CUDA_CHECK(cudaGetLastError());
            std::vector<u32> G(size_t(k_n) * k_n);
            CUDA_CHECK(cudaMemcpy(G.data(), d_G.p, G.size() * 4, cudaMemcpyDeviceToHost));
            part = greedy_partition(G, k_n);
        } else {
            part = basic_partition(k_n, std::max<u32>(arity, 2));
        }
        std::vector<int> next;
        for (const std::vector<u32> &grp : part) {
            if (grp.size() == 1) { next.push_back(cur[grp[0]]); continue; }   // passed through unchanged (BRWT_builders.cpp:74-76)
            const int id = int(N.size());
            N.emplace_back();
            full.emplace_back();
            full[id].alloc(std::max<u64>(W, 1));
            CUDA_CHECK(cudaMemset(full[id].p, 0, std::max<u64>(W, 1) * 8));
            if (sampled) { sub_own.emplace_back(); sub_own[id].alloc(Ws); CUDA_CHECK(cudaMemset(sub_own[id].p, 0, Ws * 8)); }
            for (u32 g : grp) {
                const int c = cur[g];
                N[id].children.push_back(c);
                N[id].cols.insert(N[id].cols.end(), N[c].cols.begin(), N[c].cols.end());
                if (W) or_into_kernel<<<u32((W + 255) / 256), 256>>>(full[id].p, full[c].p, W);
                if (sampled) or_into_kernel<<<u32((Ws + 255) / 256), 256>>>(sub_own[id].p, sub[c], Ws);
            }
            CUDA_CHECK(cudaGetLastError());
            sub.push_back(sampled ? sub_own[id].p : full[id].p);
            next.push_back(id);
        }
        cur.swap(next);
    }
    const int root = cur[0];
    CUDA_CHECK(cudaDeviceSynchronize());
    for (size_t id = 0; id < N.size(); ++id) N[id].ones = device_popcount(full[id].p, W);
    if (relax_arity > 1) relax_tree(N, root, n_rows, relax_arity);
    if (shape_out) *shape_out = N;
    if (root_out) *root_out = root;
    // BFS flattening; bn_of[flat id] = builder node
    std::vector<int> bn_of;
    shape_to_flat(N, root, n_rows, n_cols, &t, &bn_of);
    // a child's stored vector = its index vector restricted to the ones of its parent's (generate_subindex)
    PlainFill fill = [&](const FlatTree &ft, u32 *d_plain, const std::vector<u64> &lane_plain_off) {
        DeviceBuf<u64> cnt, before;
        DeviceBuf<unsigned char> tmp;
        cnt.alloc(std::max<u64>(W, 1));
        before.alloc(std::max<u64>(W, 1));
        size_t need = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, need, cnt.p, before.p, W);
        tmp.alloc(need + 16);
        const u32 g = u32((W + 255) / 256);
        for (size_t id = 0; id < ft.nodes.size() && W; ++id) {
            const FlatNode &fn = ft.nodes[id];
            const u32 li = fn.group * 4 + fn.lane;
            if (fn.parent < 0) words64_to_32_kernel<<<g, 256>>>(full[bn_of[id]].p, W, d_plain + lane_plain_off[li]);
            if (fn.children.empty()) continue;
            const u64 *pv = full[bn_of[id]].p;
            word_popc_kernel<<<g, 256>>>(pv, W, cnt.p);
            cub::DeviceScan::ExclusiveSum(tmp.p, need, cnt.p, before.p, W);
            for (int c : fn.children) {
                const FlatNode &cn = ft.nodes[c];
                pext_compact_kernel<<<g, 256>>>(full[bn_of[c]].p, pv, before.p, W, d_plain + lane_plain_off[cn.group * 4 + cn.lane]);
            }
            CUDA_CHECK(cudaGetLastError());
        }
        CUDA_CHECK(cudaDeviceSynchronize());
    };
    return upload_brwt(std::move(t), device, 0, 1, &fill);
}

// BRWT::serialize (BRWT.cpp:113-128) from the device image: pre-order; per node its RangePartition (the stream's own
// when the matrix was parsed from one, else derived from the column arrangements: group g = positions of child g's
// columns among this node's), its vector read back from the slot arena and re-encoded as the reference's
// rrr_vector<63> (ggq_rrr_write.hpp: byte-identical arrays), the child count, the children.
static void serialize_brwt(const Matrix *m, ByteSink &sink) {
    const FlatTree &t = m->tree;
    if (m->kind != Matrix::BRWT || t.nodes.empty())
        throw std::runtime_error("serialize needs the host tree (not kept for column-major matrices and image files)");
    if (t.n_shards > 1) throw std::runtime_error("a column shard cannot serialize the whole matrix");
    CUDA_CHECK(cudaSetDevice(m->device));
    if (t.n_cols == 0) {   // BRWT(): no groups, the rrr_vector of an empty bit vector, no children
        sink.u64be(0);
        write_rrr63(sink, nullptr, 0);
        sink.u64be(0);
        return;
    }
    DeviceBuf<u32> d_plain;
    std::vector<u64> words;
    std::vector<u32> pos_of;   // column -> position in the current node's arrangement
    struct Frame { u32 id; size_t next; };
    std::vector<Frame> stack = {{0u, 0}};
    while (!stack.empty()) {
        Frame &f = stack.back();
        const FlatNode &n = t.nodes[f.id];
        if (f.next == 0) {
            // RangePartition
            if (!n.partition.empty()) {
                sink.u64be(n.partition.size());
                for (const auto &g : n.partition) write_u32_vector(sink, g);
            } else if (n.children.empty()) {
                sink.u64be(1);
                std::vector<u32> ident(n.cols.size());
                for (size_t i = 0; i < ident.size(); ++i) ident[i] = u32(i);
                write_u32_vector(sink, ident);
            } else {
                pos_of.assign(t.n_cols, 0);
                for (size_t i = 0; i < n.cols.size(); ++i) pos_of[n.cols[i]] = u32(i);
                sink.u64be(n.children.size());
                std::vector<u32> grp;
                for (int c : n.children) {
                    grp.clear();
                    for (u32 col : t.nodes[c].cols) grp.push_back(pos_of[col]);
                    write_u32_vector(sink, grp);
                }
            }
            // nonzero_rows_
            const u32 li = n.group * 4 + n.lane;
            const u64 hot = t.group_hot[n.group], nb = t.group_nblocks[n.group];
            const u64 n_words = (n.size + 63) / 64;
            const u64 dw = slot_data_words(t.slot_bits);   // data words per block
            words.assign(std::max<u64>(n_words, (dw * nb + 1) / 2 + 1), 0);
            if (nb) {
                if (d_plain.n < dw * nb) d_plain.alloc(dw * nb + dw * nb / 8);
                if (m->wide()) unpack_lane_kernel<true><<<u32((nb + 255) / 256), 256>>>(m->slots.p, hot, nb, li & 3, d_plain.p);
                else unpack_lane_kernel<false><<<u32((nb + 255) / 256), 256>>>(m->slots.p, hot, nb, li & 3, d_plain.p);
                CUDA_CHECK(cudaGetLastError());
                CUDA_CHECK(cudaMemcpy(words.data(), d_plain.p, size_t(dw) * nb * 4, cudaMemcpyDeviceToHost));   // LE u32 pairs = u64
            }
            write_rrr63(sink, reinterpret_cast<const uint64_t *>(words.data()), n.size);
            sink.u64be(n.children.size());
        }
        if (f.next < n.children.size()) {
            const u32 c = u32(n.children[f.next++]);
            stack.push_back({c, 0});
        } else {
            stack.pop_back();
        }
    }
}

// The same builder for the IMPLICIT matrix of the benchmark (SURVEY.md 8d generator), at sizes where the columns
// cannot be materialised (C4: 1.2 TB): the greedy pairing only ever looks at the <= 10^6 sampled rows, so those are
// generated on the device (sample_implicit_kernel); the popcounts relax needs come from a count-only pass of K6 over
// ALL rows on the binary tree; the final shape is then built by K6 as usual.  Same tree as build_brwt_columns on the
// materialised columns (tests/test_builder_gpu.py).
__global__ void sample_implicit_kernel(const u64 *idx, u64 n_idx, u32 col, u64 n_cols, u64 seed_mul, u64 threshold, u64 *sub) {
    const u64 w = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (w * 64 >= n_idx) return;
    u64 x = 0;
    for (u32 b = 0; b < 64 && w * 64 + b < n_idx; ++b)
        if (mix64(seed_mul + idx[w * 64 + b] * n_cols + col) < threshold) x |= 1ull << b;
    sub[w] = x;
}

static Matrix *synthesize_brwt_greedy(u64 n_rows, u32 n_cols, u64 threshold, u64 seed, u64 relax_arity, int device) {
    CUDA_CHECK(cudaSetDevice(device));
    if (n_cols == 0 || n_rows == 0) return synthesize_brwt(n_rows, n_cols, threshold, seed, 2, device);
    const u64 seed_mul = seed * 0x9E3779B97F4A7C15ULL;
    // sampled rows (all of them up to 10^6), and every column restricted to them
    std::vector<u64> idx = build_row_sample(n_rows);
    DeviceBuf<u64> d_idx;
    upload_vec(d_idx, idx);
    const u64 Ws = (idx.size() + 63) / 64;
    std::vector<BuildNode> N(n_cols);
    std::vector<DeviceBuf<u64>> sub(n_cols);
    for (u32 j = 0; j < n_cols; ++j) {
        N[j].cols = {j};
        sub[j].alloc(Ws);
        sample_implicit_kernel<<<u32((Ws + 255) / 256), 256>>>(d_idx.p, idx.size(), j, n_cols, seed_mul, threshold, sub[j].p);
    }
    CUDA_CHECK(cudaGetLastError());
    std::vector<int> cur(n_cols);
    for (u32 j = 0; j < n_cols; ++j) cur[j] = int(j);
    DeviceBuf<u32> d_G;
    DeviceBuf<const u64 *> d_ptrs;
    while (cur.size() > 1) {
        const u32 k_n = u32(cur.size());
        std::vector<const u64 *> ptrs(k_n);
        for (u32 i = 0; i < k_n; ++i) ptrs[i] = sub[cur[i]].p;
        d_ptrs.alloc(k_n);
        CUDA_CHECK(cudaMemcpy(d_ptrs.p, ptrs.data(), size_t(k_n) * sizeof(const u64 *), cudaMemcpyHostToDevice));
        d_G.alloc(size_t(k_n) * k_n);
        const u32 tiles = (k_n + kGramTile - 1) / kGramTile;
        gram_kernel<<<dim3(tiles, tiles), 256>>>(d_ptrs.p, k_n, Ws, d_G.p);
        CUDA_CHECK(cudaGetLastError());
        std::vector<u32> G(size_t(k_n) * k_n);
        CUDA_CHECK(cudaMemcpy(G.data(), d_G.p, G.size() * 4, cudaMemcpyDeviceToHost));
        std::vector<int> next;
        for (const std::vector<u32> &grp : greedy_partition(G, k_n)) {
            if (grp.size() == 1) { next.push_back(cur[grp[0]]); continue; }
            const int id = int(N.size());
            N.emplace_back();
            sub.emplace_back();
            sub[id].alloc(Ws);
            CUDA_CHECK(cudaMemset(sub[id].p, 0, Ws * 8));
            for (u32 g : grp) {
                const int c = cur[g];
                N[id].children.push_back(c);
                N[id].cols.insert(N[id].cols.end(), N[c].cols.begin(), N[c].cols.end());
                or_into_kernel<<<u32((Ws + 255) / 256), 256>>>(sub[id].p, sub[c].p, Ws);
            }
            CUDA_CHECK(cudaGetLastError());
            next.push_back(id);
        }
        cur.swap(next);
    }
    const int root = cur[0];
    CUDA_CHECK(cudaDeviceSynchronize());
    sub.clear();
    d_G.release();
    std::vector<int> bn_of;
    if (relax_arity > 1) {
        // exact popcounts of every node's index vector over ALL rows: K6's count pass on the binary tree
        FlatTree probe;
        shape_to_flat(N, root, n_rows, n_cols, &probe, &bn_of);
        std::vector<u64> ones;
        synthesize_brwt_shape(std::move(probe), n_rows, n_cols, threshold, seed, device, 0, 1, &ones);
        for (size_t id = 0; id < bn_of.size(); ++id) N[bn_of[id]].ones = ones[id];
        relax_tree(N, root, n_rows, relax_arity);
    }
    FlatTree t;
    shape_to_flat(N, root, n_rows, n_cols, &t, &bn_of);
    for (FlatNode &fn : t.nodes) { fn.ones = 0; if (fn.parent >= 0) fn.size = 0; }   // K6 fills them in
    return synthesize_brwt_shape(std::move(t), n_rows, n_cols, threshold, seed, device);
}

// ------------------------------------------------------------------------------------------
// Workspaces
// ------------------------------------------------------------------------------------------
Workspace::Workspace(const Matrix *m_) : m(m_) {
    CUDA_CHECK(cudaSetDevice(m->device));
    CUDA_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    if (const char *e = std::getenv("GGQ_L2_PERSIST")) {
        // Experiment (north star: "node headers staged in ... the L2 persisting window"): pin the descent's group
        // records in the persisting part of L2.  Measured on C4: no effect — the 22 KB table is hit
        // every iteration by every warp and never leaves L1/L2 anyway — so it stays off by default.
        if (*e && *e != '0' && m->kind == Matrix::BRWT && (m->group_rec16.p || m->group_rec.p)) {
            cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, size_t(4) << 20);
            cudaStreamAttrValue attr{};
            attr.accessPolicyWindow.base_ptr = m->compact_rec ? (void *)m->group_rec16.p : (void *)m->group_rec.p;
            attr.accessPolicyWindow.num_bytes = m->compact_rec ? m->group_rec16.bytes() : m->group_rec.bytes();
            attr.accessPolicyWindow.hitRatio = 1.0f;
            attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
            attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
            CUDA_CHECK(cudaStreamSetAttribute(stream, cudaStreamAttributeAccessPolicyWindow, &attr));
        }
    }
    d_ctrl.alloc(16);   // 8 control words of K1 + the completion ticket of the direct label kernel
    d_err.alloc(1);
    CUDA_CHECK(cudaMallocHost(&h_ctrl, 32 * sizeof(u64)));
    CUDA_CHECK(cudaEventCreate(&ev0));
    CUDA_CHECK(cudaEventCreate(&ev1));
    CUDA_CHECK(cudaEventCreate(&ev2));
    CUDA_CHECK(cudaEventCreate(&ev3));
    CUDA_CHECK(cudaEventCreateWithFlags(&ev_ctrl, cudaEventDisableTiming));
}

Workspace::~Workspace() {
    cudaSetDevice(m->device);
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    if (ev2) cudaEventDestroy(ev2);
    if (ev3) cudaEventDestroy(ev3);
    if (ev_ctrl) cudaEventDestroy(ev_ctrl);
    if (stream) cudaStreamDestroy(stream);
    if (h_ctrl) cudaFreeHost(h_ctrl);
    if (h_stage) cudaFreeHost(h_stage);
}

void Workspace::reserve_rows(u64 n) {
    if (n + 1 <= cap_rows) return;
    u64 c = std::max<u64>(n + 1, cap_rows * 2);
    d_rows.alloc(c);
    d_row_tmp_off.alloc(c);
    d_row_cnt.alloc(c);
    d_offsets.alloc(c);
    d_bytes8.alloc(c);
    if (m->kind == Matrix::BRWT && m->sort_bits) {
        d_iota.alloc(c);
        d_perm.alloc(c);
        d_rows_sorted.alloc(c);
        iota_kernel<<<u32((c + 255) / 256), 256, 0, stream>>>(d_iota.p, c);
        size_t sneed = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, sneed, d_rows.p, d_rows_sorted.p, d_iota.p, d_perm.p, c, 0, 64, stream);
        d_sort_tmp.alloc(sneed + 16);
        sort_tmp_bytes = sneed;
    }
    d_blk.alloc(std::max<u64>((c + kScanBlock - 1) / kScanBlock + 1, 64));
    size_t need = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, need, d_row_cnt.p, d_offsets.p, c, stream);
    d_cub.alloc(need + 16);
    cub_bytes = need;
    cap_rows = c;
}

void Workspace::stage_reserve(size_t bytes) {
    if (bytes <= h_stage_bytes) return;
    if (h_stage) cudaFreeHost(h_stage);
    h_stage = nullptr;
    size_t nb = std::max(bytes, h_stage_bytes * 2);
    CUDA_CHECK(cudaMallocHost(&h_stage, nb));
    h_stage_bytes = nb;
}

Workspace *Matrix::acquire() const {
    std::lock_guard<std::mutex> lk(pool_mu);
    if (!pool.empty()) {
        Workspace *w = pool.back();
        pool.pop_back();
        return w;
    }
    return new Workspace(this);
}
void Matrix::release(Workspace *w) const {
    w->pipelined = false;
    std::lock_guard<std::mutex> lk(pool_mu);
    pool.push_back(w);
}
Matrix::~Matrix() {
    for (Workspace *w : pool) delete w;
    delete cm;
}

// ------------------------------------------------------------------------------------------
// get_rows on the device: rows already in ws.d_rows or caller memory
// ------------------------------------------------------------------------------------------
// K1 launch on the workspace stream (d_ctrl must have been cleared)
// GGQ_TRACE=1: device timeline of the pipelined host API (events on the chunk streams, printed to stderr at the end
// of the call).  Debug aid for the chunk plan; off by default, no cost when off.
struct Trace {
    struct Mark { cudaEvent_t ev; const char *what; u64 chunk; };
    std::vector<Mark> marks;
    void mark(cudaStream_t st, const char *what, u64 chunk) {
        cudaEvent_t e;
        CUDA_CHECK(cudaEventCreate(&e));
        CUDA_CHECK(cudaEventRecord(e, st));
        marks.push_back({e, what, chunk});
    }
    void dump() {
        if (marks.empty()) return;
        CUDA_CHECK(cudaDeviceSynchronize());
        for (const Mark &k : marks) {
            float ms = 0;
            cudaEventElapsedTime(&ms, marks[0].ev, k.ev);
            std::fprintf(stderr, "[ggq trace] chunk %3llu %-10s %9.3f ms\n", (unsigned long long)k.chunk, k.what, ms);
            if (&k != &marks[0]) cudaEventDestroy(k.ev);
        }
        cudaEventDestroy(marks[0].ev);
        marks.clear();
    }
};
static thread_local Trace *g_trace = nullptr;
static thread_local u64 g_trace_chunk = 0;
#define GGQ_MARK(st, what) do { if (g_trace) g_trace->mark(st, what, g_trace_chunk); } while (0)

static void launch_rows_kernel(const Matrix *m, Workspace *ws, const u64 *proc_rows, const u32 *perm, u64 n) {
    RowsArgs a;
    a.rows = proc_rows; a.perm = perm; a.n = n;
    a.tmp_cols = ws->d_tmp.p; a.tmp_cap = ws->d_tmp.n;
    a.cursor = ws->d_ctrl.p;
    a.row_tmp_off = ws->d_row_tmp_off.p; a.row_cnt = ws->d_row_cnt.p;
    const bool safe = m->stack_safe.load(std::memory_order_relaxed);
    const u32 full_grid = ws->pipelined ? m->rows_grid_p : m->rows_grid;
    const u32 cap_spill = safe ? m->cap_safe : m->cap_spill;   // safe: spill instantiation with the provable capacity
    u32 grid_cap = m->rows_grid;
    if (cap_spill) {   // at most ~1 GiB of spill area per workspace: very wide trees run the safe mode on fewer CTAs
        grid_cap = u32(std::max<u64>(1, std::min<u64>(grid_cap, (u64(1) << 27) / (u64(kRowsWarps) * cap_spill))));
        const size_t need = size_t(grid_cap) * kRowsWarps * cap_spill;
        if (ws->d_spill.n < need) { CUDA_CHECK(cudaStreamSynchronize(ws->stream)); ws->d_spill.alloc(need); }
    }
    a.spill = ws->d_spill.p; a.cap_s = m->cap_s; a.cap_spill = cap_spill;
    a.bm_words = m->bm_words; a.rb = m->rb;
    a.sum_words = rows_sum_words(m->bm_words); a.sum_magic = u32((u64(1) << 32) / a.sum_words) + 1u;
    a.err = reinterpret_cast<int *>(ws->d_ctrl.p + 4);
    a.root_child = m->root_child; a.root_sl = m->root_sl; a.root_multi_sb = m->root_multi_sb;
    a.root_base = m->root_base; a.any_more = m->any_more; a.root_all_ones = m->root_all_ones;
    u64 batches = (n + m->rb - 1) / m->rb;
    u32 grid = u32(std::min<u64>(std::min(full_grid, grid_cap), (batches + kRowsWarps - 1) / kRowsWarps));
    ws->ran_safe = safe;
    ws->ctrl_clean = false;   // (K1 writes the control words; only the direct label kernel leaves them zeroed again)
    Img img = m->img;
    void *kargs[] = {&img, &a};
    CUDA_CHECK(cudaLaunchKernel(rows_kernel_fn(safe || m->cap_spill != 0, m->img.multi_leaf != 0, m->img.group_rec16 != nullptr, m->wide()), dim3(grid),
                                dim3(kRowsWarps * 32), kargs, ws->pipelined ? m->rows_smem_p : m->rows_smem, ws->stream));
}

// Enqueue (no synchronisation) everything of a row query up to the point where the host must learn the result
// size: batch ordering, descent, offset scan, control words to pinned memory.
static void brwt_rows_begin(const Matrix *m, Workspace *ws, const u64 *d_rows, u64 n) {
    ws->reserve_rows(n);
    cudaStream_t st = ws->stream;
    ws->pending_sort_launches = 0;
    if (n == 0) {
        CUDA_CHECK(cudaMemsetAsync(ws->d_offsets.p, 0, 8, st));
        return;
    }
    if (ws->d_tmp.n == 0) ws->d_tmp.alloc(std::max<u64>(n * 16, 1 << 16));
    // order the batch by row id (top sort_bits bits): monotone walks through every node vector
    const u64 *proc_rows = d_rows;
    const u32 *perm = nullptr;
    if (m->sort_bits && n >= m->sort_min_n && ws->pipelined && m->sort_bits <= 16 && n < (u64(1) << 32)) {
        // counting sort (histogram, block scan of the bucket counters, scatter): small kernels that fit beside the
        // previous chunk's descent
        const u32 n_buckets = 1u << m->sort_bits, shift = m->row_bits - m->sort_bits;
        if (ws->d_order.n < 2 * size_t(n_buckets) + 16) ws->d_order.alloc(2 * size_t(n_buckets) + 16);
        u32 *cnt = ws->d_order.p, *cursor = ws->d_order.p + n_buckets;   // counts | exclusive offsets (n_buckets + 1)
        const u32 g = u32(std::min<u64>((n + 1023) / 1024, u64(m->sm_count) * 8));
        const u32 nbk = (n_buckets + 1 + kScanBlock - 1) / kScanBlock;
        if (ws->d_blk.n < nbk) ws->d_blk.alloc(nbk + 16);
        CUDA_CHECK(cudaMemsetAsync(cnt, 0, size_t(n_buckets) * 4, st));
        order_hist_kernel<<<g, 256, 0, st>>>(d_rows, n, shift, n_buckets - 1, cnt);
        cnt_block_sum_kernel<<<nbk, 256, 0, st>>>(cnt, n_buckets, ws->d_blk.p);
        block_sum_scan_kernel<<<1, 256, 0, st>>>(ws->d_blk.p, nbk);
        cnt_block_scan_kernel<u32><<<nbk, 256, 0, st>>>(cnt, n_buckets, ws->d_blk.p, cursor);
        order_scatter_kernel<<<g, 256, 0, st>>>(d_rows, n, shift, n_buckets - 1, cursor, ws->d_rows_sorted.p, ws->d_perm.p);
        CUDA_CHECK(cudaGetLastError());
        proc_rows = ws->d_rows_sorted.p;
        perm = ws->d_perm.p;
        ws->pending_sort_launches = 5;
    } else if (m->sort_bits && n >= m->sort_min_n) {
        size_t sb = ws->sort_tmp_bytes;
        int end_bit = int(m->row_bits), begin_bit = int(m->row_bits - m->sort_bits);
        cub::DeviceRadixSort::SortPairs(ws->d_sort_tmp.p, sb, d_rows, ws->d_rows_sorted.p, ws->d_iota.p, ws->d_perm.p, n,
                                        begin_bit, end_bit, st);
        CUDA_CHECK(cudaGetLastError());
        proc_rows = ws->d_rows_sorted.p;
        perm = ws->d_perm.p;
        ws->pending_sort_launches = 1 + (end_bit - begin_bit + 7) / 8;   // cub onesweep: histogram + one kernel per 8-bit digit
    }
    CUDA_CHECK(cudaMemsetAsync(ws->d_ctrl.p, 0, 8 * sizeof(u64), st));   // incl. the overflow flag in ctrl[4]
    GGQ_MARK(st, "sorted");
    CUDA_CHECK(cudaEventRecord(ws->ev0, st));
    launch_rows_kernel(m, ws, proc_rows, perm, n);
    CUDA_CHECK(cudaEventRecord(ws->ev1, st));
    GGQ_MARK(st, "descent");
    if (ws->pipelined) {
        // the host learns the chunk's size as soon as the descent is done; offsets by the small scan kernels
        publish_ctrl_kernel<<<1, 32, 0, st>>>(ws->d_ctrl.p, ws->h_ctrl);
        CUDA_CHECK(cudaEventRecord(ws->ev_ctrl, st));
        if (ws->hist_mode) { GGQ_MARK(st, "scan+ctrl"); return; }   // count_labels: no CSR, no offsets
        const u32 nb = u32((n + 1 + kScanBlock - 1) / kScanBlock);
        if (ws->d_blk.n < nb) throw std::runtime_error("scan scratch not reserved");
        cnt_block_sum_kernel<<<nb, 256, 0, st>>>(ws->d_row_cnt.p, n, ws->d_blk.p);
        block_sum_scan_kernel<<<1, 256, 0, st>>>(ws->d_blk.p, nb);
        cnt_block_scan_kernel<u64><<<nb, 256, 0, st>>>(ws->d_row_cnt.p, n, ws->d_blk.p, ws->d_offsets.p);
        CUDA_CHECK(cudaGetLastError());
    } else {
        size_t cb = ws->cub_bytes;
        // offsets[0..n] : exclusive sum over n+1 counts (the last one is a zero pad)
        CUDA_CHECK(cudaMemsetAsync(ws->d_row_cnt.p + n, 0, 4, st));
        cub::DeviceScan::ExclusiveSum(ws->d_cub.p, cb, ws->d_row_cnt.p, ws->d_offsets.p, n + 1, st);
        CUDA_CHECK(cudaMemcpyAsync(ws->h_ctrl, ws->d_ctrl.p, 8 * sizeof(u64), cudaMemcpyDeviceToHost, st));
    }
    GGQ_MARK(st, "scan+ctrl");
}

// Wait for brwt_rows_begin, then assemble the CSR (enqueued, not awaited).  The chunk arena grows and the query
// reruns if it was too small (steady state never does).
static void brwt_rows_finish(const Matrix *m, Workspace *ws, const u64 *d_rows, u64 n, int col_bytes, u64 *total_out) {
    cudaStream_t st = ws->stream;
    if (n == 0) {
        CUDA_CHECK(cudaStreamSynchronize(st));
        *total_out = 0;
        ws->launches = 0;
        return;
    }
    for (int attempt = 0; attempt < 3; ++attempt) {
        if (ws->pipelined) CUDA_CHECK(cudaEventSynchronize(ws->ev_ctrl));
        else CUDA_CHECK(cudaStreamSynchronize(st));
        if (*reinterpret_cast<int *>(ws->h_ctrl + 4)) {
            // the estimated stack capacity was too small for this batch: nothing was written out of bounds (K1 checks
            // before it pushes); switch the matrix to the provable capacity and redo the batch
            if (ws->ran_safe) throw std::runtime_error("descent stack overflow beyond the provable bound (engine bug)");
            m->stack_safe.store(true, std::memory_order_relaxed);
            brwt_rows_begin(m, ws, d_rows, n);
            continue;
        }
        u64 total = ws->h_ctrl[0];
        ws->last_probes = ws->h_ctrl[2];
        ws->last_touches = ws->h_ctrl[3];
        ws->last_code_probes = ws->h_ctrl[5];
        ws->last_code_escapes = ws->h_ctrl[6];
        ws->kernel_timed = true;
        // descent + ordering + (2 cub scan kernels | pipelined: ctrl publish + 3 scan kernels, none for a histogram)
        ws->launches = 1 + ws->pending_sort_launches + (ws->pipelined ? (ws->hist_mode ? 1 : 4) : 2);
        if (total > ws->d_tmp.n) {  // chunk arena too small: grow and redo
            ws->d_tmp.alloc(total + total / 8 + 1024);
            brwt_rows_begin(m, ws, d_rows, n);
            continue;
        }
        *total_out = total;
        if (ws->hist_mode) return;   // ids stay in the chunk arena (order is irrelevant to a histogram)
        size_t out_bytes = size_t(total) * col_bytes;
        if (ws->d_out.n < out_bytes) ws->d_out.alloc(out_bytes + out_bytes / 8 + 1024);
        if (total) {
            u32 g = u32(((n + kGatherRows - 1) / kGatherRows * 8 + 255) / 256);
            if (col_bytes == 8)
                gather_rows_kernel<u64><<<g, 256, 0, st>>>(ws->d_tmp.p, ws->d_row_tmp_off.p, ws->d_offsets.p, n,
                                                          reinterpret_cast<u64 *>(ws->d_out.p));
            else if (col_bytes == 2)
                gather_rows_kernel<unsigned short><<<g, 256, 0, st>>>(ws->d_tmp.p, ws->d_row_tmp_off.p, ws->d_offsets.p, n,
                                                                     reinterpret_cast<unsigned short *>(ws->d_out.p));
            else
                gather_rows_kernel<u32><<<g, 256, 0, st>>>(ws->d_tmp.p, ws->d_row_tmp_off.p, ws->d_offsets.p, n,
                                                          reinterpret_cast<u32 *>(ws->d_out.p));
            CUDA_CHECK(cudaGetLastError());
            ws->launches += 1;
        }
        return;
    }
    throw std::runtime_error("chunk arena did not converge");
}

static void brwt_rows_device(const Matrix *m, Workspace *ws, const u64 *d_rows, u64 n, int col_bytes, u64 *total_out) {
    brwt_rows_begin(m, ws, d_rows, n);
    brwt_rows_finish(m, ws, d_rows, n, col_bytes, total_out);
}

void rows_device(const Matrix *m, Workspace *ws, const u64 *d_rows, u64 n, int col_bytes, u64 *total_out) {
    CUDA_CHECK(cudaSetDevice(m->device));
    if (m->kind == Matrix::BRWT) brwt_rows_device(m, ws, d_rows, n, col_bytes, total_out);
    else colmajor_rows_device(m, ws, d_rows, n, col_bytes, total_out);
    m->last_probes = ws->last_probes;
}

struct RowsResult {
    const Matrix *m;
    Workspace *ws;   // owns the pinned staging the pointers below refer to
    u64 n, total;
    const u64 *offsets, *cols;
};

}  // namespace ggq

// ==========================================================================================
// C ABI
// ==========================================================================================
using namespace ggq;

struct ggq_matrix : ggq::Matrix {};
struct ggq_workspace : ggq::Workspace { using ggq::Workspace::Workspace; };
struct ggq_rows_result : ggq::RowsResult {};
struct ggq_comm : ggq::Comm {};

#define GGQ_TRY try {
#define GGQ_CATCH                                                                   \
    } catch (const ggq::ParseError &e) { return fail(GGQ_ERR_FORMAT, e.what());     \
    } catch (const ggq::LimitError &e) { return fail(GGQ_ERR_LIMIT, e.what());      \
    } catch (const ggq::CudaError &e) { return fail(GGQ_ERR_CUDA, e.what());        \
    } catch (const ggq::NcclError &e) { return fail(GGQ_ERR_NCCL, e.what());        \
    } catch (const std::bad_alloc &) { return fail(GGQ_ERR_OOM, "out of memory");   \
    } catch (const std::exception &e) { return fail(GGQ_ERR_INTERNAL, e.what()); }

extern "C" {

const char *ggq_last_error(void) { return g_last_error.c_str(); }

int ggq_device_count(int *count) {
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e != cudaSuccess) { *count = 0; return fail(GGQ_ERR_CUDA, cudaGetErrorString(e)); }
    *count = c;
    return GGQ_OK;
}

static int brwt_load_reader(Reader &in, int device, ggq_matrix **out, u32 shard = 0, u32 n_shards = 1) {
    GGQ_TRY
    if (!out) return fail(GGQ_ERR_ARG, "null out");
    ParsedNode root;
    const char *eager = std::getenv("GGQ_LOAD_EAGER");   // (A/B of the load path's host memory: scripts/load_probe.py)
    load_brwt_node(in, &root, /*lazy=*/!(eager && *eager && *eager != '0'));   // vectors stay undecoded until their upload
    FlatTree t;
    flatten_parsed(std::move(root), &t);
    if (n_shards < 1 || shard >= n_shards) return fail(GGQ_ERR_ARG, "bad shard");
    *out = static_cast<ggq_matrix *>(upload_brwt(std::move(t), device, shard, n_shards));
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_brwt_load(ggq_read_fn read, void *ctx, int device, ggq_matrix **out) {
    if (!read) return fail(GGQ_ERR_ARG, "null reader");
    Reader in(read, ctx);
    return brwt_load_reader(in, device, out);
}

int ggq_brwt_load_buffer(const uint8_t *buf, size_t len, size_t *consumed, int device, ggq_matrix **out) {
    if (!buf && len) return fail(GGQ_ERR_ARG, "null buffer");
    Reader in(buf, len);
    int rc = brwt_load_reader(in, device, out);
    if (consumed) *consumed = in.consumed();
    return rc;
}

int ggq_brwt_load_shard(ggq_read_fn read, void *ctx, int device, uint32_t shard, uint32_t n_shards, ggq_matrix **out) {
    if (!read) return fail(GGQ_ERR_ARG, "null reader");
    Reader in(read, ctx);
    return brwt_load_reader(in, device, out, shard, n_shards);
}

int ggq_brwt_load_shard_buffer(const uint8_t *buf, size_t len, size_t *consumed, int device, uint32_t shard,
                               uint32_t n_shards, ggq_matrix **out) {
    if (!buf && len) return fail(GGQ_ERR_ARG, "null buffer");
    Reader in(buf, len);
    int rc = brwt_load_reader(in, device, out, shard, n_shards);
    if (consumed) *consumed = in.consumed();
    return rc;
}

int ggq_brwt_synthesize_shard(uint64_t n_rows, uint64_t n_cols, uint64_t threshold, uint64_t seed, uint32_t arity,
                              int device, uint32_t shard, uint32_t n_shards, ggq_matrix **out) {
    GGQ_TRY
    if (!out || arity < 2 || n_shards < 1 || shard >= n_shards) return fail(GGQ_ERR_ARG, "bad argument");
    if (n_cols >= (1ull << 31)) return fail(GGQ_ERR_LIMIT, "too many columns");
    *out = static_cast<ggq_matrix *>(synthesize_brwt(n_rows, u32(n_cols), threshold, seed, arity, device, shard, n_shards));
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_brwt_shard_plan_buffer(const uint8_t *buf, size_t len, uint32_t n_shards, uint32_t *col_owner, uint32_t *cut_depth) {
    GGQ_TRY
    if ((!buf && len) || n_shards < 1) return fail(GGQ_ERR_ARG, "bad argument");
    Reader in(buf, len);
    ParsedNode root;
    load_brwt_node(in, &root);
    FlatTree t;
    flatten_parsed(std::move(root), &t);
    assign_groups(&t);
    plan_shards(&t, 0, n_shards);
    if (col_owner)
        for (u64 c = 0; c < t.n_cols; ++c) col_owner[c] = t.col_owner[c];
    if (cut_depth) *cut_depth = t.cut_depth;
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_shard_info(const ggq_matrix *m, uint32_t *shard, uint32_t *n_shards, uint8_t *col_owned) {
    if (!m) return fail(GGQ_ERR_ARG, "null matrix");
    if (shard) *shard = m->tree.shard;
    if (n_shards) *n_shards = m->tree.n_shards;
    if (col_owned)
        for (u64 c = 0; c < m->tree.n_cols; ++c) col_owned[c] = m->tree.col_owned.empty() ? 1 : m->tree.col_owned[c];
    return GGQ_OK;
}

int ggq_colmajor_load(ggq_read_fn read, void *ctx, int device, ggq_matrix **out) {
    if (!read) return fail(GGQ_ERR_ARG, "null reader");
    GGQ_TRY
    Reader in(read, ctx);
    *out = static_cast<ggq_matrix *>(colmajor_load(in, device));
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_colmajor_load_buffer(const uint8_t *buf, size_t len, size_t *consumed, int device, ggq_matrix **out) {
    if (!buf && len) return fail(GGQ_ERR_ARG, "null buffer");
    GGQ_TRY
    Reader in(buf, len);
    *out = static_cast<ggq_matrix *>(colmajor_load(in, device));
    if (consumed) *consumed = in.consumed();
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_brwt_synthesize(uint64_t n_rows, uint64_t n_cols, uint64_t threshold, uint64_t seed, uint32_t arity,
                        int device, ggq_matrix **out) {
    GGQ_TRY
    if (!out || arity < 2) return fail(GGQ_ERR_ARG, "bad argument");
    if (n_cols >= (1ull << 31)) return fail(GGQ_ERR_LIMIT, "too many columns");
    *out = static_cast<ggq_matrix *>(synthesize_brwt(n_rows, u32(n_cols), threshold, seed, arity, device));
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_brwt_build_columns(const uint64_t *col_words, uint64_t n_rows, uint64_t n_cols, int partitioner, uint32_t arity,
                           uint64_t relax_arity, int device, ggq_matrix **out) {
    GGQ_TRY
    if (!out || (!col_words && n_rows && n_cols) || partitioner < 0 || partitioner > 1 || (partitioner == 0 && arity < 2))
        return fail(GGQ_ERR_ARG, "bad argument");
    if (n_rows >= kMaxRows || n_cols >= (u64(1) << 31)) return fail(GGQ_ERR_LIMIT, "matrix exceeds engine limits");
    *out = static_cast<ggq_matrix *>(build_brwt_columns(reinterpret_cast<const u64 *>(col_words), n_rows, u32(n_cols),
                                                        partitioner, arity, relax_arity, device));
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_brwt_synthesize_greedy(uint64_t n_rows, uint64_t n_cols, uint64_t threshold, uint64_t seed, uint64_t relax_arity,
                               int device, ggq_matrix **out) {
    GGQ_TRY
    if (!out) return fail(GGQ_ERR_ARG, "null out");
    if (n_rows >= kMaxRows || n_cols >= (u64(1) << 31)) return fail(GGQ_ERR_LIMIT, "matrix exceeds engine limits");
    *out = static_cast<ggq_matrix *>(synthesize_brwt_greedy(n_rows, u32(n_cols), threshold, seed, relax_arity, device));
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_tree_dfs_columns(const ggq_matrix *m, uint32_t *dfs_cols) {
    if (!m || (!dfs_cols && m->tree.n_cols)) return fail(GGQ_ERR_ARG, "bad argument");
    if (m->kind != Matrix::BRWT) return fail(GGQ_ERR_ARG, "not a BRWT");
    for (size_t i = 0; i < m->tree.leaf_col.size(); ++i) dfs_cols[i] = m->tree.leaf_col[i];
    return GGQ_OK;
}

int ggq_tree_shape(const ggq_matrix *m, uint64_t *n_nodes, int64_t *parent, uint32_t *n_children, uint32_t *n_columns,
                   uint64_t *size, uint64_t *ones, uint32_t *first_column) {
    if (!m || !n_nodes) return fail(GGQ_ERR_ARG, "bad argument");
    if (m->kind != Matrix::BRWT || m->tree.nodes.empty()) return fail(GGQ_ERR_ARG, "no host tree (column-major matrix or image file)");
    const auto &N = m->tree.nodes;
    *n_nodes = N.size();
    for (size_t i = 0; i < N.size(); ++i) {
        if (parent) parent[i] = N[i].parent;
        if (n_children) n_children[i] = u32(N[i].children.size());
        if (n_columns) n_columns[i] = u32(N[i].cols.size());
        if (size) size[i] = N[i].size;
        if (ones) ones[i] = N[i].ones;
        if (first_column) first_column[i] = N[i].cols.empty() ? ~0u : N[i].cols[0];
    }
    return GGQ_OK;
}

int ggq_brwt_serialize(const ggq_matrix *m, ggq_write_fn write, void *ctx, uint64_t *n_bytes) {
    GGQ_TRY
    if (!m || !write) return fail(GGQ_ERR_ARG, "bad argument");
    ByteSink sink;
    sink.write = write;
    sink.ctx = ctx;
    serialize_brwt(m, sink);
    if (n_bytes) *n_bytes = sink.total;
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_brwt_serialize_buffer(const ggq_matrix *m, uint8_t **out, size_t *len) {
    GGQ_TRY
    if (!m || !out || !len) return fail(GGQ_ERR_ARG, "bad argument");
    std::string bytes;
    ByteSink sink;
    sink.out = &bytes;
    serialize_brwt(m, sink);
    uint8_t *buf = static_cast<uint8_t *>(std::malloc(std::max<size_t>(bytes.size(), 1)));
    if (!buf) throw std::bad_alloc();
    std::memcpy(buf, bytes.data(), bytes.size());
    *out = buf;
    *len = bytes.size();
    return GGQ_OK;
    GGQ_CATCH
}

// Files: the .brwt.annodbg matrix part as the reference reads / writes it, through buffered stdio — no copy of the
// stream in host memory on either side.
int ggq_brwt_serialize_file(const ggq_matrix *m, const char *path, uint64_t *n_bytes) {
    GGQ_TRY
    if (!m || !path) return fail(GGQ_ERR_ARG, "bad argument");
    FILE *f = std::fopen(path, "wb");
    if (!f) return fail(GGQ_ERR_ARG, std::string("cannot create ") + path);
    std::vector<char> buf(size_t(8) << 20);
    std::setvbuf(f, buf.data(), _IOFBF, buf.size());
    ByteSink sink;
    sink.write = [](void *ctx, const void *src, size_t n) -> size_t { return std::fwrite(src, 1, n, static_cast<FILE *>(ctx)); };
    sink.ctx = f;
    int rc = GGQ_OK;
    try {
        serialize_brwt(m, sink);
    } catch (...) {
        std::fclose(f);
        throw;
    }
    if (std::fclose(f) != 0) rc = fail(GGQ_ERR_INTERNAL, std::string("write failed: ") + path);
    if (n_bytes) *n_bytes = sink.total;
    return rc;
    GGQ_CATCH
}

int ggq_brwt_load_file(const char *path, int device, ggq_matrix **out) {
    if (!path || !out) return fail(GGQ_ERR_ARG, "bad argument");
    FILE *f = std::fopen(path, "rb");
    if (!f) return fail(GGQ_ERR_FORMAT, std::string("cannot open ") + path);
    std::vector<char> buf(size_t(8) << 20);
    std::setvbuf(f, buf.data(), _IOFBF, buf.size());
    Reader in([](void *ctx, void *dst, size_t n) -> size_t { return std::fread(dst, 1, n, static_cast<FILE *>(ctx)); }, f);
    const int rc = brwt_load_reader(in, device, out);
    std::fclose(f);
    return rc;
}

// Host only (no device needed): the bytes rrr_vector<63>(bits).serialize() writes — the building block of
// ggq_brwt_serialize, exported so that the CPU tests can pin it against the reference byte for byte.
int ggq_rrr63_encode_buffer(const uint64_t *words, uint64_t n_bits, uint8_t **out, size_t *len) {
    GGQ_TRY
    if (!out || !len || (!words && n_bits)) return fail(GGQ_ERR_ARG, "bad argument");
    std::string bytes;
    ByteSink sink;
    sink.out = &bytes;
    std::vector<u64> padded;   // clear the padding bits of the last word
    const u64 n_words = (n_bits + 63) / 64;
    const u64 *w = reinterpret_cast<const u64 *>(words);
    if (n_bits & 63) {
        padded.assign(w, w + n_words);
        padded[n_words - 1] &= (u64(1) << (n_bits & 63)) - 1;
        w = padded.data();
    }
    write_rrr63(sink, reinterpret_cast<const uint64_t *>(w), n_bits);
    uint8_t *buf = static_cast<uint8_t *>(std::malloc(std::max<size_t>(bytes.size(), 1)));
    if (!buf) throw std::bad_alloc();
    std::memcpy(buf, bytes.data(), bytes.size());
    *out = buf;
    *len = bytes.size();
    return GGQ_OK;
    GGQ_CATCH
}

// Host only: the inverse — the loader's own decoder (read_rrr63_raw + decode_rrr63) on one serialized rrr_vector<63>.
int ggq_rrr63_decode_buffer(const uint8_t *buf, size_t len, size_t *consumed, uint64_t *words, uint64_t n_words, uint64_t *n_bits) {
    GGQ_TRY
    if ((!buf && len) || !n_bits || (!words && n_words)) return fail(GGQ_ERR_ARG, "bad argument");
    Reader in(buf, len);
    RrrRaw raw;
    read_rrr63_raw(in, &raw);
    if (consumed) *consumed = in.consumed();
    PlainBits bits;
    decode_rrr63(raw, &bits);
    *n_bits = bits.size;
    const u64 need = (bits.size + 63) / 64;
    if (need > n_words) return fail(GGQ_ERR_ARG, "capacity too small");
    for (u64 i = 0; i < need; ++i) {   // u32 words (zero padded) -> LSB-first u64 words
        const u64 lo = 2 * i < bits.w.size() ? bits.w[2 * i] : 0, hi = 2 * i + 1 < bits.w.size() ? bits.w[2 * i + 1] : 0;
        words[i] = lo | (hi << 32);
    }
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_brwt_inspect_buffer(const uint8_t *buf, size_t len, size_t *consumed, uint64_t stats[8]) {
    GGQ_TRY
    if ((!buf && len) || !stats) return fail(GGQ_ERR_ARG, "bad argument");
    Reader in(buf, len);
    ParsedNode root;
    load_brwt_node(in, &root);
    if (consumed) *consumed = in.consumed();
    FlatTree t;
    flatten_parsed(std::move(root), &t);
    t.slot_bits = default_slot_bits();   // (the image size below is that of the layout a load would use now)
    assign_groups(&t);
    assign_geometry(&t);
    u64 rel = 0, tcs = 0;
    for (const FlatNode &n : t.nodes) {
        tcs += n.size;
        if (n.children.empty()) rel += n.ones * n.cols.size();
    }
    stats[0] = t.n_rows; stats[1] = t.n_cols; stats[2] = t.nodes.size(); stats[3] = rel; stats[4] = tcs;
    stats[5] = t.n_groups(); stats[6] = t.total_slots * 16 + t.total_sel * 4 + t.total_sb * 8; stats[7] = t.max_depth;
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_brwt_inspect_columns(const uint8_t *buf, size_t len, uint32_t *dfs_cols, uint32_t *depth) {
    GGQ_TRY
    if (!buf && len) return fail(GGQ_ERR_ARG, "bad argument");
    Reader in(buf, len);
    ParsedNode root;
    load_brwt_node(in, &root);
    FlatTree t;
    flatten_parsed(std::move(root), &t);
    assign_groups(&t);
    if (dfs_cols)
        for (size_t i = 0; i < t.leaf_col.size(); ++i) dfs_cols[i] = t.leaf_col[i];
    if (depth)
        for (u64 c = 0; c < t.n_cols; ++c) depth[c] = t.col_path_off[c + 1] - t.col_path_off[c] - 1;
    return GGQ_OK;
    GGQ_CATCH
}

void ggq_free(ggq_matrix *m) {
    if (!m) return;
    cudaSetDevice(m->device);
    delete static_cast<ggq::Matrix *>(m);
}

uint64_t ggq_num_rows(const ggq_matrix *m) { return m->tree.n_rows; }
uint64_t ggq_num_columns(const ggq_matrix *m) { return m->tree.n_cols; }
uint64_t ggq_num_relations(const ggq_matrix *m) { return m->num_relations; }
uint64_t ggq_num_nodes(const ggq_matrix *m) { return m->n_nodes; }
uint64_t ggq_total_column_size(const ggq_matrix *m) { return m->total_column_size; }
uint64_t ggq_device_bytes(const ggq_matrix *m) { return m->device_bytes; }
uint32_t ggq_slot_bits(const ggq_matrix *m) { return m && m->kind == Matrix::BRWT ? m->tree.slot_bits : 0; }

int ggq_set_default_option(const char *name, uint64_t value) {
    GGQ_TRY
    if (!name) return fail(GGQ_ERR_ARG, "bad argument");
    if (std::string(name) == "wide_slots") { g_default_wide.store(value ? 1 : 0, std::memory_order_relaxed); return GGQ_OK; }
    return fail(GGQ_ERR_ARG, "unknown option");
    GGQ_CATCH
}
uint64_t ggq_last_probe_count(const ggq_matrix *m) { return m->last_probes; }
int ggq_leaf_code_info(const ggq_matrix *m, uint64_t what[4]) {
    if (!m || !what) return fail(GGQ_ERR_ARG, "bad argument");
    what[0] = m->n_coded; what[1] = m->code_bytes; what[2] = uint64_t(m->code_escape_rate * 1e6 + 0.5);
    what[3] = m->use_codes && m->n_coded;
    return GGQ_OK;
}

int ggq_set_option(ggq_matrix *m, const char *name, uint64_t value) {
    GGQ_TRY
    if (!m || !name) return fail(GGQ_ERR_ARG, "bad argument");
    if (m->kind != Matrix::BRWT) return GGQ_OK;
    std::lock_guard<std::mutex> lk(m->pool_mu);
    for (Workspace *w : m->pool) delete w;   // scratch sizes depend on the configuration
    m->pool.clear();
    CUDA_CHECK(cudaSetDevice(m->device));
    std::string k(name);
    if (k == "leaf_codes") {   // 0: plain descent everywhere (A/B, and the reference's touch count V_t); 1: use the codes
        m->use_codes = value != 0;
        CUDA_CHECK(cudaSetDevice(m->device));
        CUDA_CHECK(cudaDeviceSynchronize());
        m->upload_codes_tables();
        return GGQ_OK;
    }
    if (k == "rows_per_warp") m->configure_rows(u32(value), m->sort_bits);
    else if (k == "sort_bits") m->configure_rows(m->rb, u32(std::min<uint64_t>(value, 64)));
    else if (k == "sort_min_rows") m->sort_min_n = value;
    else return fail(GGQ_ERR_ARG, "unknown option");
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_workspace_create(const ggq_matrix *m, ggq_workspace **out) {
    GGQ_TRY
    *out = new ggq_workspace(m);
    return GGQ_OK;
    GGQ_CATCH
}
void ggq_workspace_free(ggq_workspace *ws) { delete ws; }
void *ggq_workspace_stream(ggq_workspace *ws) { return ws->stream; }
double ggq_workspace_stat(const ggq_workspace *ws, int what) {
    switch (what) {
        case GGQ_STAT_KERNEL_MS: {
            float ms = 0.f;
            if (!ws->kernel_timed || cudaEventElapsedTime(&ms, ws->ev0, ws->ev1) != cudaSuccess) return -1.0;
            return ms;
        }
        case GGQ_STAT_PROBES: return double(ws->last_probes);
        case GGQ_STAT_TOUCHES: return double(ws->last_touches);
        case GGQ_STAT_LAUNCHES: return double(ws->launches);
        case GGQ_STAT_EXCHANGE_MS: {
            float ms = 0.f;
            if (!ws->exchange_timed || cudaEventSynchronize(ws->ev3) != cudaSuccess ||
                cudaEventElapsedTime(&ms, ws->ev2, ws->ev3) != cudaSuccess) return -1.0;
            return ms;
        }
        case GGQ_STAT_EXCHANGE_BYTES: return double(ws->exchange_bytes);
        case GGQ_STAT_CODE_PROBES: return double(ws->last_code_probes);
        case GGQ_STAT_CODE_ESCAPES: return double(ws->last_code_escapes);
        default: return -1.0;
    }
}

int ggq_get_rows_device(const ggq_matrix *m, ggq_workspace *ws, const uint64_t *d_rows, uint64_t n, int col_bytes,
                        const uint64_t **d_offsets, const void **d_cols, uint64_t *n_cols) {
    GGQ_TRY
    if (!m || !ws || (!d_rows && n) || (col_bytes != 4 && col_bytes != 8)) return fail(GGQ_ERR_ARG, "bad argument");
    u64 total = 0;
    rows_device(m, ws, reinterpret_cast<const u64 *>(d_rows), n, col_bytes, &total);
    if (d_offsets) *d_offsets = reinterpret_cast<const uint64_t *>(ws->d_offsets.p);
    if (d_cols) *d_cols = ws->d_out.p;
    if (n_cols) *n_cols = total;
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_merge_shards_device(ggq_workspace *ws, uint64_t n, int n_shards, const uint64_t *const *d_offsets,
                            const uint32_t *const *d_cols, int col_bytes, const uint64_t **d_out_offsets,
                            const void **d_out_cols, uint64_t *n_cols) {
    GGQ_TRY
    if (!ws || n_shards < 1 || n_shards > kMaxShards || !d_offsets || !d_cols || (col_bytes != 4 && col_bytes != 8))
        return fail(GGQ_ERR_ARG, "bad argument");
    CUDA_CHECK(cudaSetDevice(ws->m->device));
    ShardPtrs p;
    p.n = n_shards;
    for (int r = 0; r < n_shards; ++r) { p.off[r] = reinterpret_cast<const u64 *>(d_offsets[r]); p.cols[r] = d_cols[r]; }
    u64 total = 0;
    merge_shards(ws, p, n, col_bytes, &total);
    if (d_out_offsets) *d_out_offsets = reinterpret_cast<const uint64_t *>(ws->m_off.p);
    if (d_out_cols) *d_out_cols = ws->m_out.p;
    if (n_cols) *n_cols = total;
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_image_save(const ggq_matrix *m, const char *path) {
    GGQ_TRY
    if (!m || !path) return fail(GGQ_ERR_ARG, "bad argument");
    image_save(m, path);
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_image_load(const char *path, int device, ggq_matrix **out) {
    GGQ_TRY
    if (!path || !out) return fail(GGQ_ERR_ARG, "bad argument");
    *out = static_cast<ggq_matrix *>(image_load(path, device));
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_comm_unique_id(uint8_t id[128]) {
    GGQ_TRY
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    ncclUniqueId u;
    NCCL_CHECK(NcclApi::get().GetUniqueId(&u));
    std::memcpy(id, &u, 128);
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_comm_init(const uint8_t id[128], int rank, int n_ranks, int device, ggq_comm **out) {
    GGQ_TRY
    if (!id || !out || rank < 0 || rank >= n_ranks) return fail(GGQ_ERR_ARG, "bad argument");
    CUDA_CHECK(cudaSetDevice(device));
    std::unique_ptr<ggq_comm> c(new ggq_comm());
    ncclUniqueId u;
    std::memcpy(&u, id, 128);
    NCCL_CHECK(NcclApi::get().CommInitRank(&c->comm, n_ranks, u, rank));
    c->rank = rank; c->n_ranks = n_ranks; c->device = device;
    *out = c.release();
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_comm_enable_peer(ggq_comm *c, ggq_workspace *ws, uint64_t max_rows, uint64_t max_ids) {
    GGQ_TRY
    if (!c || !ws || !max_rows) return fail(GGQ_ERR_ARG, "bad argument");
    if (c->n_ranks > kMaxShards) return fail(GGQ_ERR_LIMIT, "too many shards");
    CUDA_CHECK(cudaSetDevice(c->device));
    const int R = c->n_ranks;
    c->close_peer();   // a second call re-pins and re-maps (collective, like the first)
    // pin the partial-CSR buffers at the stated capacity (they must not be reallocated while mapped by peers)
    ws->reserve_rows(max_rows);
    const size_t out_bytes = size_t(std::max<u64>(max_ids, 1)) * 4;
    if (ws->d_out.n < out_bytes) ws->d_out.alloc(out_bytes);
    if (ws->g_tot.n < size_t(kMaxShards) + 1) ws->g_tot.alloc(kMaxShards + 1);
    struct Handles { cudaIpcMemHandle_t off, cols; };
    static_assert(sizeof(Handles) == 128, "cudaIpcMemHandle_t size");
    Handles mine;
    CUDA_CHECK(cudaIpcGetMemHandle(&mine.off, ws->d_offsets.p));
    CUDA_CHECK(cudaIpcGetMemHandle(&mine.cols, ws->d_out.p));
    DeviceBuf<unsigned char> d_send, d_recv;
    d_send.alloc(sizeof(Handles));
    d_recv.alloc(sizeof(Handles) * size_t(R));
    CUDA_CHECK(cudaMemcpyAsync(d_send.p, &mine, sizeof(Handles), cudaMemcpyHostToDevice, ws->stream));
    NCCL_CHECK(NcclApi::get().AllGather(d_send.p, d_recv.p, sizeof(Handles), ncclUint8, c->comm, ws->stream));
    std::vector<Handles> all(static_cast<size_t>(R));
    CUDA_CHECK(cudaMemcpyAsync(all.data(), d_recv.p, sizeof(Handles) * size_t(R), cudaMemcpyDeviceToHost, ws->stream));
    CUDA_CHECK(cudaStreamSynchronize(ws->stream));
    for (int r = 0; r < R; ++r) {
        if (r == c->rank) continue;
        void *po = nullptr, *pc = nullptr;
        CUDA_CHECK(cudaIpcOpenMemHandle(&po, all[size_t(r)].off, cudaIpcMemLazyEnablePeerAccess));
        c->peer_off[r] = static_cast<const u64 *>(po);
        CUDA_CHECK(cudaIpcOpenMemHandle(&pc, all[size_t(r)].cols, cudaIpcMemLazyEnablePeerAccess));
        c->peer_cols[r] = static_cast<const u32 *>(pc);
    }
    c->peer_ws = ws;
    c->peer_max_rows = max_rows;
    c->peer_max_ids = max_ids;
    c->peer = true;
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_comm_peer_active(const ggq_comm *c) { return c && c->peer ? 1 : 0; }

void ggq_comm_free(ggq_comm *c) {
    if (!c) return;
    if (c->comm) NcclApi::get().CommDestroy(c->comm);
    delete c;
}

static int rows_sharded_c(const ggq_matrix *m, ggq_workspace *ws, ggq_comm *comm, const uint64_t *d_rows, uint64_t n,
                          int col_bytes, bool slice, uint64_t *first, uint64_t *count, const uint64_t **d_offsets,
                          const void **d_cols, uint64_t *n_cols) {
    GGQ_TRY
    if (!m || !ws || !comm || (!d_rows && n) || (col_bytes != 4 && col_bytes != 8)) return fail(GGQ_ERR_ARG, "bad argument");
    if (m->kind != Matrix::BRWT) return fail(GGQ_ERR_ARG, "not a BRWT");
    u64 total = 0, q0 = 0, nq = n;
    rows_sharded_device(m, ws, comm, reinterpret_cast<const u64 *>(d_rows), n, col_bytes, &total, slice, &q0, &nq);
    if (first) *first = q0;
    if (count) *count = nq;
    if (d_offsets) *d_offsets = reinterpret_cast<const uint64_t *>(ws->m_off.p);
    if (d_cols) *d_cols = ws->m_out.p;
    if (n_cols) *n_cols = total;
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_get_rows_sharded_device(const ggq_matrix *m, ggq_workspace *ws, ggq_comm *comm, const uint64_t *d_rows,
                                uint64_t n, int col_bytes, const uint64_t **d_offsets, const void **d_cols,
                                uint64_t *n_cols) {
    return rows_sharded_c(m, ws, comm, d_rows, n, col_bytes, false, nullptr, nullptr, d_offsets, d_cols, n_cols);
}

int ggq_get_rows_sharded_slice_device(const ggq_matrix *m, ggq_workspace *ws, ggq_comm *comm, const uint64_t *d_rows,
                                      uint64_t n, int col_bytes, uint64_t *first_query, uint64_t *n_queries,
                                      const uint64_t **d_offsets, const void **d_cols, uint64_t *n_cols) {
    return rows_sharded_c(m, ws, comm, d_rows, n, col_bytes, true, first_query, n_queries, d_offsets, d_cols, n_cols);
}

// host rows -> device query -> host CSR in the workspace's pinned staging buffer
static void rows_host(const ggq_matrix *m, Workspace *ws, const uint64_t *rows, u64 n, u64 *total) {
    CUDA_CHECK(cudaSetDevice(m->device));
    ws->reserve_rows(n);
    if (n) CUDA_CHECK(cudaMemcpyAsync(ws->d_rows.p, rows, n * 8, cudaMemcpyHostToDevice, ws->stream));
    rows_device(m, ws, ws->d_rows.p, n, 8, total);
}

int ggq_get_rows(const ggq_matrix *m, const uint64_t *rows, uint64_t n, ggq_rows_result **out) {
    GGQ_TRY
    if (!m || !out || (!rows && n)) return fail(GGQ_ERR_ARG, "bad argument");
    Workspace *ws = m->acquire();
    try {
        u64 total = 0;
        // small batches (a scalar get_row is a batch of one): memset + descent + one finishing CTA that writes the CSR
        // straight into the pinned staging buffer, one synchronisation (small_finish_kernel)
        if (m->kind == Matrix::BRWT && n >= 1 && n <= u64(kSmallRows) && n < m->sort_min_n) {
            CUDA_CHECK(cudaSetDevice(m->device));
            const u64 cap_ids = u64(1) << 16;
            ws->reserve_rows(n);
            // pinned block: [total | status | offsets (n + 1) | ids (cap_ids) | completion flag | the rows]
            ws->stage_reserve((2 + n + 1 + cap_ids + 1 + n) * 8);
            if (ws->d_tmp.n == 0) ws->d_tmp.alloc(std::max<u64>(n * 16, 1 << 16));
            cudaStream_t st = ws->stream;
            u64 *host = reinterpret_cast<u64 *>(ws->h_stage);
            // Latency form (as in labels_batch_core): the descent reads the row ids straight from the pinned block, the
            // control words were left cleared by the previous small call, the host polls a flag — two launches in all.
            static const bool copy_form = [] { const char *e = std::getenv("GGQ_DIRECT_COPY"); return e && *e && *e != '0'; }();
            u64 *flag = host + 2 + n + 1 + cap_ids, *h_rows = flag + 1;
            const u64 *d_small_rows = ws->d_rows.p;
            if (copy_form) {
                CUDA_CHECK(cudaMemcpyAsync(ws->d_rows.p, rows, n * 8, cudaMemcpyHostToDevice, st));
            } else {
                std::memcpy(h_rows, rows, n * 8);
                *flag = 0;
                d_small_rows = h_rows;
            }
            if (!ws->ctrl_clean || copy_form) CUDA_CHECK(cudaMemsetAsync(ws->d_ctrl.p, 0, 9 * sizeof(u64), st));
            launch_rows_kernel(m, ws, d_small_rows, nullptr, n);
            small_finish_kernel<<<1, 256, 0, st>>>(ws->d_row_cnt.p, ws->d_row_tmp_off.p, ws->d_tmp.p, ws->d_tmp.n, ws->d_ctrl.p,
                                                  u32(n), host, cap_ids, copy_form ? nullptr : flag);
            CUDA_CHECK(cudaGetLastError());
            bool flagged = false;
            if (!copy_form) {   // (bounded: a failed launch never sets the flag)
                volatile u64 *f = flag;
                const auto t_end = std::chrono::steady_clock::now() + std::chrono::milliseconds(2);
                for (u32 spin = 0; !(flagged = *f != 0); ++spin)
                    if ((spin & 255u) == 255u && std::chrono::steady_clock::now() > t_end) break;
                std::atomic_thread_fence(std::memory_order_acquire);
            }
            if (!flagged) CUDA_CHECK(cudaStreamSynchronize(st));
            if (!copy_form) ws->ctrl_clean = true;
            if (host[1] & 1) m->stack_safe.store(true, std::memory_order_relaxed);   // estimate too small: general path, safe mode
            if (!(host[1] & 3)) {
                ws->launches = 2; ws->kernel_timed = false; ws->last_probes = 0; ws->last_touches = 0;
                ggq_rows_result *r = new ggq_rows_result();
                r->m = m; r->ws = ws; r->n = n; r->total = host[0];
                r->offsets = host + 2;
                r->cols = host + 2 + n + 1;
                *out = r;
                return GGQ_OK;
            }
            // rows too long for the staging buffer (or the chunk arena had to grow): the general path below
        }
        rows_host(m, ws, rows, n, &total);
        size_t off_bytes = (n + 1) * 8, col_bytes = total * 8;
        ws->stage_reserve(off_bytes + col_bytes + 8);
        CUDA_CHECK(cudaMemcpyAsync(ws->h_stage, ws->d_offsets.p, off_bytes, cudaMemcpyDeviceToHost, ws->stream));
        if (total)
            CUDA_CHECK(cudaMemcpyAsync(ws->h_stage + off_bytes, ws->d_out.p, col_bytes, cudaMemcpyDeviceToHost, ws->stream));
        CUDA_CHECK(cudaStreamSynchronize(ws->stream));
        ggq_rows_result *r = new ggq_rows_result();
        r->m = m; r->ws = ws; r->n = n; r->total = total;
        r->offsets = reinterpret_cast<const u64 *>(ws->h_stage);
        r->cols = reinterpret_cast<const u64 *>(ws->h_stage + off_bytes);
        *out = r;
    } catch (...) {
        m->release(ws);
        throw;
    }
    return GGQ_OK;
    GGQ_CATCH
}

const uint64_t *ggq_rows_offsets(const ggq_rows_result *r) { return reinterpret_cast<const uint64_t *>(r->offsets); }
const uint64_t *ggq_rows_columns(const ggq_rows_result *r) { return reinterpret_cast<const uint64_t *>(r->cols); }
uint64_t ggq_rows_count(const ggq_rows_result *r) { return r->n; }
void ggq_rows_result_free(ggq_rows_result *r) {
    if (!r) return;
    r->m->release(r->ws);
    delete r;
}

// Large batches are cut into chunks that alternate between two workspaces (streams): while chunk k's CSR
// travels to the host, chunk k+1 is sorted and descended — the PCIe copy, not the kernel, bounds this call.
// rows per chunk of the pipelined host calls (GGQ_CHUNK_LOG2 overrides for experiments)
static u64 pipeline_chunk() {
    const char *e = std::getenv("GGQ_CHUNK_LOG2");
    const u64 l = e && *e ? std::strtoull(e, nullptr, 10) : 21;
    return u64(1) << std::min<u64>(std::max<u64>(l, 10), 30);
}
// count_labels returns only num_columns counters, so nothing but the first host->device copy is exposed: larger
// chunks (measured on C4: 2^20 -> 1.03e9, 2^22 -> 1.12e9 rows/s); get_rows is best at 2^20 (device->host tail)
static const u64 kCountChunk = u64(1) << 22;
static const u64 kPipelineDepth = 4;   // workspaces (streams) a pipelined host call rotates over

// rows: u64 ids (row_bytes 8) or u32 ids (4).  Row boundaries come back as CSR offsets (count_bytes 0: n + 1 u64 in
// `offsets`) or as per-row counts of count_bytes = 1 / 2 / 4 bytes in `counts_v`.  Column ids of col_bytes = 8 / 4 / 2.
static int get_rows_into(const ggq_matrix *m, const void *rows_v, int row_bytes, uint64_t n, uint64_t *offsets,
                         void *counts_v, int count_bytes, void *cols_v, uint64_t cols_capacity, uint64_t *n_cols,
                         int col_bytes) {
    unsigned char *cols = static_cast<unsigned char *>(cols_v);
    unsigned char *counts = static_cast<unsigned char *>(counts_v);
    const unsigned char *rows = static_cast<const unsigned char *>(rows_v);
    GGQ_TRY
    if (!m || (!offsets && !count_bytes) || (count_bytes && !counts && n) || (!rows && n)) return fail(GGQ_ERR_ARG, "bad argument");
    if (col_bytes == 2 && m->tree.n_cols > 65536) return fail(GGQ_ERR_ARG, "u16 column ids need a matrix of at most 65536 columns");
    if (row_bytes == 4 && m->tree.n_rows > (u64(1) << 32)) return fail(GGQ_ERR_ARG, "u32 row ids need a matrix of at most 2^32 rows");
    CUDA_CHECK(cudaSetDevice(m->device));
    // Chunks rotate over kPipelineDepth workspaces (streams).  Chunk k+1 is enqueued (H2D, ordering, descent, scan)
    // BEFORE the host waits for chunk k's size, so the device never idles on a host round trip, and the device->host
    // copy of chunk k overlaps the descent of the chunks after it.
    // Chunk plan: full chunks in the middle (large launches cost less per row), geometric ramps at both
    // ends — the first device->host copy starts early and the last one (which nothing overlaps) is short.
    std::vector<u64> start = {0};
    {
        const u64 c = pipeline_chunk();
        std::vector<u64> ramp;
        for (u64 r = std::max<u64>(c / 8, u64(1) << 18); r < c; r *= 2) ramp.push_back(r);
        u64 ramp_sum = 0;
        for (u64 r : ramp) ramp_sum += r;
        if (!ramp.empty() && n >= 2 * ramp_sum + c && !std::getenv("GGQ_NO_RAMP")) {
            for (u64 r : ramp) start.push_back(start.back() + r);
            while (n - start.back() > ramp_sum + c) start.push_back(start.back() + c);
            const u64 rest = n - start.back() - ramp_sum;   // in (0, c]
            start.push_back(start.back() + rest);
            for (size_t i = ramp.size(); i-- > 1;) start.push_back(start.back() + ramp[i]);
            start.push_back(n);
        } else {
            const u64 cs = std::getenv("GGQ_NO_RAMP") ? c : std::min<u64>(c, u64(1) << 20);
            while (n - start.back() > cs) start.push_back(start.back() + cs);
            if (n > start.back() || start.size() == 1) start.push_back(n);
        }
    }
    const u64 n_chunks = start.size() - 1;
    const bool brwt = m->kind == Matrix::BRWT;
    std::vector<Workspace *> w(size_t(std::min<u64>(n_chunks, kPipelineDepth)), nullptr);
    int rc = GGQ_OK;
    if (std::getenv("GGQ_TRACE") && n_chunks > 1 && !g_trace) g_trace = new Trace();
    try {
        const bool room = n_chunks > 1 && !std::getenv("GGQ_NO_PIPE_ROOM");
        for (auto &p : w) {
            p = m->acquire();
            p->pipelined = room;
            if (count_bytes) CUDA_CHECK(cudaMemsetAsync(p->d_err.p, 0, sizeof(int), p->stream));
        }
        u64 base = 0;
        bool fits = true;
        int launches = 0;
        u64 probes = 0, touches = 0;
        auto chunk_rows = [&](u64 k) { return start[k + 1] - start[k]; };
        auto begin = [&](u64 k) {
            Workspace *ws = w[k % w.size()];
            const u64 nc = chunk_rows(k);
            ws->reserve_rows(nc);
            g_trace_chunk = k;
            GGQ_MARK(ws->stream, "begin");
            if (nc && row_bytes == 8) {
                CUDA_CHECK(cudaMemcpyAsync(ws->d_rows.p, rows + start[k] * 8, nc * 8, cudaMemcpyHostToDevice, ws->stream));
            } else if (nc) {
                if (ws->d_rows32.n < nc) ws->d_rows32.alloc(ws->cap_rows);
                CUDA_CHECK(cudaMemcpyAsync(ws->d_rows32.p, rows + start[k] * 4, nc * 4, cudaMemcpyHostToDevice, ws->stream));
                widen_rows_kernel<<<u32((nc + 255) / 256), 256, 0, ws->stream>>>(ws->d_rows32.p, nc, ws->d_rows.p);
                CUDA_CHECK(cudaGetLastError());
            }
            GGQ_MARK(ws->stream, "h2d");
            if (brwt) brwt_rows_begin(m, ws, ws->d_rows.p, nc);
        };
        begin(0);
        for (u64 k = 0; k < n_chunks; ++k) {
            if (k + 1 < n_chunks) begin(k + 1);
            Workspace *ws = w[k % w.size()];
            const u64 nc = chunk_rows(k), r0 = start[k];
            u64 total = 0;
            g_trace_chunk = k;
            if (brwt) brwt_rows_finish(m, ws, ws->d_rows.p, nc, col_bytes, &total);
            else rows_device(m, ws, ws->d_rows.p, nc, col_bytes, &total);
            GGQ_MARK(ws->stream, "gather");
            launches += ws->launches; probes += ws->last_probes; touches += ws->last_touches;
            if (count_bytes && nc) {   // per-row counts, narrowed on the device (colmajor: from the offsets)
                if (ws->d_cnt_out.n < size_t(nc) * 4) ws->d_cnt_out.alloc(size_t(ws->cap_rows) * 4);
                if (!brwt) {
                    offsets_to_counts_kernel<<<u32((nc + 255) / 256), 256, 0, ws->stream>>>(ws->d_offsets.p, nc, ws->d_row_cnt.p);
                }
                u32 *flag = reinterpret_cast<u32 *>(ws->d_err.p);   // sticky over the chunks of this call
                const u32 g = u32((nc + 255) / 256);
                if (count_bytes == 1) narrow_counts_kernel<unsigned char><<<g, 256, 0, ws->stream>>>(ws->d_row_cnt.p, nc, ws->d_cnt_out.p, flag);
                else if (count_bytes == 2) narrow_counts_kernel<unsigned short><<<g, 256, 0, ws->stream>>>(ws->d_row_cnt.p, nc, reinterpret_cast<unsigned short *>(ws->d_cnt_out.p), flag);
                else narrow_counts_kernel<u32><<<g, 256, 0, ws->stream>>>(ws->d_row_cnt.p, nc, reinterpret_cast<u32 *>(ws->d_cnt_out.p), flag);
                CUDA_CHECK(cudaGetLastError());
                ++launches;
                CUDA_CHECK(cudaMemcpyAsync(counts + r0 * count_bytes, ws->d_cnt_out.p, nc * count_bytes, cudaMemcpyDeviceToHost, ws->stream));
            } else if (!count_bytes) {
                if (base) {
                    add_base_kernel<<<u32((nc + 1 + 255) / 256), 256, 0, ws->stream>>>(ws->d_offsets.p, nc + 1, base);
                    CUDA_CHECK(cudaGetLastError());
                    ++launches;
                }
                CUDA_CHECK(cudaMemcpyAsync(offsets + r0, ws->d_offsets.p, (nc + 1) * 8, cudaMemcpyDeviceToHost, ws->stream));
            }
            if (base + total > cols_capacity || (total && !cols)) fits = false;
            if (fits && total)
                CUDA_CHECK(cudaMemcpyAsync(cols + base * col_bytes, ws->d_out.p, total * col_bytes, cudaMemcpyDeviceToHost, ws->stream));
            GGQ_MARK(ws->stream, "d2h");
            base += total;
        }
        if (count_bytes)
            for (Workspace *ws : w)
                CUDA_CHECK(cudaMemcpyAsync(ws->h_ctrl + 16, ws->d_err.p, sizeof(int), cudaMemcpyDeviceToHost, ws->stream));
        for (Workspace *ws : w)
            if (ws) CUDA_CHECK(cudaStreamSynchronize(ws->stream));
        bool count_overflow = false;
        if (count_bytes)
            for (Workspace *ws : w) count_overflow |= *reinterpret_cast<const int *>(ws->h_ctrl + 16) != 0;
        if (g_trace) { g_trace->dump(); delete g_trace; g_trace = nullptr; }
        w[0]->launches = launches; w[0]->last_probes = probes; w[0]->last_touches = touches;
        m->last_probes = probes;
        if (n_cols) *n_cols = base;
        if (!fits) rc = fail(GGQ_ERR_ARG, "cols_capacity too small");
        else if (count_overflow) rc = fail(GGQ_ERR_LIMIT, "a row has more columns than count_bytes can hold (counts are saturated)");
    } catch (...) {
        for (Workspace *ws : w)
            if (ws) { cudaStreamSynchronize(ws->stream); m->release(ws); }
        delete g_trace; g_trace = nullptr;
        throw;
    }
    for (Workspace *ws : w)
        if (ws) m->release(ws);
    return rc;
    GGQ_CATCH
}

int ggq_get_rows_into(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t *offsets, uint64_t *cols,
                      uint64_t cols_capacity, uint64_t *n_cols) {
    return get_rows_into(m, rows, 8, n, offsets, nullptr, 0, cols, cols_capacity, n_cols, 8);
}

int ggq_get_rows_compact(const ggq_matrix *m, const void *rows, int row_bytes, uint64_t n, void *counts, int count_bytes,
                         void *cols, int col_bytes, uint64_t cols_capacity, uint64_t *n_cols) {
    if ((row_bytes != 4 && row_bytes != 8) || (count_bytes != 1 && count_bytes != 2 && count_bytes != 4) ||
        (col_bytes != 2 && col_bytes != 4 && col_bytes != 8))
        return fail(GGQ_ERR_ARG, "row_bytes must be 4 or 8, count_bytes 1, 2 or 4, col_bytes 2, 4 or 8");
    return get_rows_into(m, rows, row_bytes, n, nullptr, counts, count_bytes, cols, cols_capacity, n_cols, col_bytes);
}

int ggq_get_rows_into_u32(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t *offsets, uint32_t *cols,
                          uint64_t cols_capacity, uint64_t *n_cols) {
    return get_rows_into(m, rows, 8, n, offsets, nullptr, 0, cols, cols_capacity, n_cols, 4);
}

static void launch_col_hist(const ggq_matrix *m, Workspace *ws, const u32 *ids, u64 total, u64 *d_counts) {
    const u64 nc = std::max<u64>(m->tree.n_cols, 1);
    const u32 smem_cols = 12288;   // 48 KB of u32 counters
    const size_t smem = nc <= smem_cols ? nc * 4 : 0;
    const u32 grid = u32(std::min<u64>(u64(m->sm_count) * 4, (total + 511) / 512));
    col_hist_kernel<<<grid, 512, smem, ws->stream>>>(ids, total, u32(nc), smem_cols, d_counts);
    CUDA_CHECK(cudaGetLastError());
    ws->launches += 1;
}

// device part of count_labels: counts (u64 x num_columns) in ws->d_counts
static void count_labels_device(const ggq_matrix *m, Workspace *ws, const u64 *d_rows, u64 n) {
    const u64 nc = std::max<u64>(m->tree.n_cols, 1);
    if (ws->d_counts.n < nc) ws->d_counts.alloc(nc);
    CUDA_CHECK(cudaMemsetAsync(ws->d_counts.p, 0, nc * 8, ws->stream));
    u64 total = 0;
    const u32 *ids = nullptr;
    if (m->kind == Matrix::BRWT) {
        ws->hist_mode = true;
        try { rows_device(m, ws, d_rows, n, 4, &total); } catch (...) { ws->hist_mode = false; throw; }
        ws->hist_mode = false;
        ids = ws->d_tmp.p;
    } else {
        rows_device(m, ws, d_rows, n, 4, &total);
        ids = reinterpret_cast<const u32 *>(ws->d_out.p);
    }
    if (total) launch_col_hist(m, ws, ids, total, ws->d_counts.p);
}

// Large BRWT batches: chunks rotate over kPipelineDepth streams like get_rows_into — the host->device copy of a
// chunk and its ordering overlap the descent of the previous one; every chunk's histogram kernel adds into the
// counters of the first workspace.
static void count_labels_pipelined(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t *counts) {
    const u64 n_chunks = (n + kCountChunk - 1) / kCountChunk;
    const u64 nc = std::max<u64>(m->tree.n_cols, 1);
    std::vector<Workspace *> w(size_t(std::min<u64>(n_chunks, kPipelineDepth)), nullptr);
    cudaEvent_t zeroed = nullptr;
    try {
        const bool room = !std::getenv("GGQ_NO_PIPE_ROOM");
        for (auto &p : w) { p = m->acquire(); p->pipelined = room; }
        CUDA_CHECK(cudaEventCreateWithFlags(&zeroed, cudaEventDisableTiming));
        if (w[0]->d_counts.n < nc) w[0]->d_counts.alloc(nc);
        CUDA_CHECK(cudaMemsetAsync(w[0]->d_counts.p, 0, nc * 8, w[0]->stream));
        CUDA_CHECK(cudaEventRecord(zeroed, w[0]->stream));
        for (size_t i = 1; i < w.size(); ++i) CUDA_CHECK(cudaStreamWaitEvent(w[i]->stream, zeroed, 0));
        auto chunk_rows = [&](u64 k) { return std::min<u64>(kCountChunk, n - k * kCountChunk); };
        auto begin = [&](u64 k) {
            Workspace *ws = w[k % w.size()];
            const u64 nr = chunk_rows(k);
            ws->reserve_rows(nr);
            CUDA_CHECK(cudaMemcpyAsync(ws->d_rows.p, rows + k * kCountChunk, nr * 8, cudaMemcpyHostToDevice, ws->stream));
            ws->hist_mode = true;
            try { brwt_rows_begin(m, ws, ws->d_rows.p, nr); } catch (...) { ws->hist_mode = false; throw; }
            ws->hist_mode = false;
        };
        int launches = 0;
        begin(0);
        for (u64 k = 0; k < n_chunks; ++k) {
            if (k + 1 < n_chunks) begin(k + 1);
            Workspace *ws = w[k % w.size()];
            u64 total = 0;
            ws->hist_mode = true;
            try { brwt_rows_finish(m, ws, ws->d_rows.p, chunk_rows(k), 4, &total); } catch (...) { ws->hist_mode = false; throw; }
            ws->hist_mode = false;
            if (total) launch_col_hist(m, ws, ws->d_tmp.p, total, w[0]->d_counts.p);
            launches += ws->launches;
        }
        for (Workspace *ws : w) CUDA_CHECK(cudaStreamSynchronize(ws->stream));
        w[0]->launches = launches;
        if (m->tree.n_cols) {
            CUDA_CHECK(cudaMemcpyAsync(counts, w[0]->d_counts.p, m->tree.n_cols * 8, cudaMemcpyDeviceToHost, w[0]->stream));
            CUDA_CHECK(cudaStreamSynchronize(w[0]->stream));
        }
    } catch (...) {
        for (Workspace *ws : w)
            if (ws) { cudaStreamSynchronize(ws->stream); m->release(ws); }
        if (zeroed) cudaEventDestroy(zeroed);
        throw;
    }
    for (Workspace *ws : w) m->release(ws);
    cudaEventDestroy(zeroed);
}

int ggq_count_labels(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t *counts) {
    GGQ_TRY
    if (!m || (!rows && n) || !counts) return fail(GGQ_ERR_ARG, "bad argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    if (m->kind == Matrix::BRWT && n > kCountChunk) {
        count_labels_pipelined(m, rows, n, counts);
        return GGQ_OK;
    }
    Workspace *ws = m->acquire();
    try {
        ws->reserve_rows(n);
        if (n) CUDA_CHECK(cudaMemcpyAsync(ws->d_rows.p, rows, n * 8, cudaMemcpyHostToDevice, ws->stream));
        count_labels_device(m, ws, ws->d_rows.p, n);
        if (m->tree.n_cols)
            CUDA_CHECK(cudaMemcpyAsync(counts, ws->d_counts.p, m->tree.n_cols * 8, cudaMemcpyDeviceToHost, ws->stream));
        CUDA_CHECK(cudaStreamSynchronize(ws->stream));
    } catch (...) {
        m->release(ws);
        throw;
    }
    m->release(ws);
    return GGQ_OK;
    GGQ_CATCH
}

// counts -> (column, count) pairs with count != 0 && count >= min_count, ascending columns; top != 0: stable sort by
// count, descending (ties keep ascending columns).  Returns the number of pairs, left in d_lab_cols/d_lab_cnt
// (or d_lab_cols2/d_lab_cnt2 after a sort: *sorted).
static u64 filter_labels_device(const ggq_matrix *m, Workspace *ws, u64 min_count, bool top, bool *sorted) {
    const u64 nc = std::max<u64>(m->tree.n_cols, 1);
    if (ws->d_lab_cols.n < nc) {
        ws->d_lab_cols.alloc(nc); ws->d_lab_cnt.alloc(nc); ws->d_lab_cols2.alloc(nc); ws->d_lab_cnt2.alloc(nc);
        ws->d_lab_n.alloc(1);
    }
    label_filter_kernel<<<1, 1024, 0, ws->stream>>>(ws->d_counts.p, u32(m->tree.n_cols), min_count, ws->d_lab_cols.p,
                                                    ws->d_lab_cnt.p, ws->d_lab_n.p);
    CUDA_CHECK(cudaGetLastError());
    ws->launches += 1;
    CUDA_CHECK(cudaMemcpyAsync(ws->h_ctrl + 8, ws->d_lab_n.p, 8, cudaMemcpyDeviceToHost, ws->stream));
    CUDA_CHECK(cudaStreamSynchronize(ws->stream));
    const u64 kept = ws->h_ctrl[8];
    *sorted = false;
    if (top && kept > 1) {
        size_t need = 0;
        cub::DeviceRadixSort::SortPairsDescending(nullptr, need, ws->d_lab_cnt.p, ws->d_lab_cnt2.p, ws->d_lab_cols.p,
                                                  ws->d_lab_cols2.p, kept, 0, 64, ws->stream);
        if (ws->d_lab_sort.n < need + 16) ws->d_lab_sort.alloc(need + 16);
        cub::DeviceRadixSort::SortPairsDescending(ws->d_lab_sort.p, need, ws->d_lab_cnt.p, ws->d_lab_cnt2.p, ws->d_lab_cols.p,
                                                  ws->d_lab_cols2.p, kept, 0, 64, ws->stream);
        CUDA_CHECK(cudaGetLastError());
        *sorted = true;
    }
    return kept;
}

static int labels_batch_core(const ggq_matrix *m, const uint64_t *rows, const uint64_t *query_offsets, uint64_t n_queries,
                             const u64 *min_counts, uint64_t *label_offsets, uint64_t *cols, uint64_t *counts,
                             uint64_t capacity, uint64_t *n_out);

static int labels_query(const ggq_matrix *m, const uint64_t *rows, uint64_t n, u64 min_count, bool top, u64 limit,
                        uint64_t *cols, uint64_t *counts, uint64_t *n_out) {
    if (!m || (!rows && n) || !n_out) return fail(GGQ_ERR_ARG, "bad argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    Workspace *ws = m->acquire();
    try {
        // one query sequence, threshold only: the direct form of the label batch with a single query — one
        // host-to-device copy, the descent, one CTA that builds the histogram from the descent's chunks, thresholds it
        // and writes the kept pairs into pinned host memory, ONE synchronisation
        const u64 nc_s = m->tree.n_cols;
        if (m->kind == Matrix::BRWT && !top && n >= 1 && n <= u64(kSmallRows) && nc_s >= 1 && nc_s <= (u64(1) << 15)) {
            m->release(ws);   // (the batch call takes its own workspace)
            const uint64_t qoff[2] = {0, n};
            uint64_t loff[2] = {0, 0};
            if (cols) return labels_batch_core(m, rows, qoff, 1, &min_count, loff, cols, counts, nc_s, n_out);
            std::vector<uint64_t> tmp(nc_s);
            return labels_batch_core(m, rows, qoff, 1, &min_count, loff, tmp.data(), counts, nc_s, n_out);
        }
        ws->reserve_rows(n);
        if (n) CUDA_CHECK(cudaMemcpyAsync(ws->d_rows.p, rows, n * 8, cudaMemcpyHostToDevice, ws->stream));
        count_labels_device(m, ws, ws->d_rows.p, n);
        bool sorted = false;
        u64 kept = m->tree.n_cols ? filter_labels_device(m, ws, min_count, top, &sorted) : 0;
        kept = std::min(kept, limit);
        if (kept && cols)
            CUDA_CHECK(cudaMemcpyAsync(cols, sorted ? ws->d_lab_cols2.p : ws->d_lab_cols.p, kept * 8, cudaMemcpyDeviceToHost, ws->stream));
        if (kept && counts)
            CUDA_CHECK(cudaMemcpyAsync(counts, sorted ? ws->d_lab_cnt2.p : ws->d_lab_cnt.p, kept * 8, cudaMemcpyDeviceToHost, ws->stream));
        CUDA_CHECK(cudaStreamSynchronize(ws->stream));
        *n_out = kept;
    } catch (...) {
        m->release(ws);
        throw;
    }
    m->release(ws);
    return GGQ_OK;
}

int ggq_get_labels(const ggq_matrix *m, const uint64_t *rows, uint64_t n, double presence_ratio, uint64_t *cols,
                   uint64_t *counts, uint64_t *n_out) {
    GGQ_TRY
    if (!(presence_ratio >= 0.0 && presence_ratio <= 1.0)) return fail(GGQ_ERR_ARG, "presence_ratio must be in [0, 1]");
    // annotate_static.cpp:79-82: size_t min_labels_discovered = ratio == 0 ? 1 : std::ceil(indices.size() * ratio)
    const u64 min_count = presence_ratio == 0 ? 1 : u64(std::ceil(double(n) * presence_ratio));
    return labels_query(m, rows, n, min_count, false, ~u64(0), cols, counts, n_out);
    GGQ_CATCH
}

int ggq_get_top_labels(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t num_top, uint64_t *cols,
                       uint64_t *counts, uint64_t *n_out) {
    GGQ_TRY
    return labels_query(m, rows, n, 1, true, num_top, cols, counts, n_out);
    GGQ_CATCH
}

// min_labels_discovered of StaticBinRelAnnotator::get_labels (annotate_static.cpp:79-81), per query
static u64 min_label_count(u64 n_indices, double ratio) {
    return ratio == 0 ? 1 : u64(std::ceil(double(n_indices) * ratio));
}

constexpr u64 kBatchSmallRows = u64(1) << 14;   // direct form of the label batch: indices of all queries together
constexpr u64 kBatchSmallQueries = 1024;
constexpr u64 kZeroCopyRows = 4096;             // direct form: up to this many indices are read straight from pinned memory

// min_counts[q]: a label of query q is kept iff its count is non-zero and >= min_counts[q]
static int labels_batch_core(const ggq_matrix *m, const uint64_t *rows, const uint64_t *query_offsets, uint64_t n_queries,
                             const u64 *min_counts, uint64_t *label_offsets, uint64_t *cols, uint64_t *counts,
                             uint64_t capacity, uint64_t *n_out) {
    GGQ_TRY
    if (!m || !query_offsets || !label_offsets || !n_out || (!min_counts && n_queries)) return fail(GGQ_ERR_ARG, "bad argument");
    if (query_offsets[0] != 0) return fail(GGQ_ERR_ARG, "query_offsets[0] must be 0");
    for (u64 q = 0; q < n_queries; ++q)
        if (query_offsets[q + 1] < query_offsets[q]) return fail(GGQ_ERR_ARG, "query_offsets must be non-decreasing");
    const u64 n = query_offsets[n_queries];
    if (!rows && n) return fail(GGQ_ERR_ARG, "bad argument");
    if (n_queries >= (u64(1) << 31)) return fail(GGQ_ERR_LIMIT, "too many queries in one batch");
    CUDA_CHECK(cudaSetDevice(m->device));
    static std::atomic<int> max_smem_cached{0};
    int max_smem = max_smem_cached.load(std::memory_order_relaxed);
    if (!max_smem) {
        CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, m->device));
        max_smem_cached.store(max_smem, std::memory_order_relaxed);
    }
    const u64 nc = m->tree.n_cols;
    const size_t hist_smem = size_t(std::max<u64>(nc, 1)) * 4;
    if (hist_smem + 1024 > size_t(max_smem)) return fail(GGQ_ERR_LIMIT, "too many columns for the per-query shared-memory histogram");
    *n_out = 0;
    if (n_queries == 0) { label_offsets[0] = 0; return GGQ_OK; }
    Workspace *ws = m->acquire();
    int rc = GGQ_OK;
    try {
        ws->reserve_rows(n);
        cudaStream_t st = ws->stream;
        if (ws->d_qoff.n < 2 * n_queries + 2) ws->d_qoff.alloc(2 * n_queries + 2);
        if (ws->d_qlab_off.n < n_queries + 2) { ws->d_qlab_off.alloc(n_queries + 2); ws->d_qkept.alloc(n_queries + 2); }
        // small batches: one pinned block [rows | query offsets | per-query minimum counts]; large ones copy the rows
        // straight from the caller's memory
        const u64 cap_pairs = u64(1) << 16;
        const bool direct = m->kind == Matrix::BRWT && n >= 1 && n <= kBatchSmallRows && n_queries <= kBatchSmallQueries;
        // direct form: [query offsets | minimum counts | rows] in one pinned block and ONE host-to-device copy into
        // d_qoff (the descent reads its rows from there)
        const size_t n_staged = direct ? size_t(n) : 0;
        const size_t in_words = (n_queries + 1) + n_queries + n_staged;
        ws->stage_reserve((in_words + 2 + 2 * n_queries + (direct ? cap_pairs : 0)) * 8);
        if (direct && ws->d_qoff.n < in_words) ws->d_qoff.alloc(std::max<size_t>(in_words, 2 * kBatchSmallQueries + 2 + kBatchSmallRows));
        u64 *h_in = reinterpret_cast<u64 *>(ws->h_stage);
        std::memcpy(h_in, query_offsets, (n_queries + 1) * 8);
        if (n_queries) std::memcpy(h_in + n_queries + 1, min_counts, n_queries * 8);
        if (n_staged) std::memcpy(h_in + 2 * n_queries + 1, rows, n * 8);
        // Latency form of the direct path (a single query sequence is launch-latency bound: every stream operation costs
        // microseconds, and concurrent callers contend for the driver): up to kZeroCopyRows indices are READ BY THE
        // KERNELS FROM THE PINNED BLOCK (each index is read once), the control words were left zeroed by the previous
        // direct call's kernel, and the host polls the completion flag in pinned memory — two launches, nothing else.
        // GGQ_DIRECT_COPY=1: the copy / memset / synchronise form (A/B).
        static const bool copy_form = [] { const char *e = std::getenv("GGQ_DIRECT_COPY"); return e && *e && *e != '0'; }();
        const bool zero_copy = direct && !copy_form && n <= kZeroCopyRows;
        const u64 *d_in = zero_copy ? h_in : ws->d_qoff.p;   // [query offsets | minimum counts | rows]
        if (!zero_copy) CUDA_CHECK(cudaMemcpyAsync(ws->d_qoff.p, h_in, in_words * 8, cudaMemcpyHostToDevice, st));
        const u64 *d_batch_rows = direct ? d_in + 2 * n_queries + 1 : ws->d_rows.p;
        if (n && !direct) CUDA_CHECK(cudaMemcpyAsync(ws->d_rows.p, rows, n * 8, cudaMemcpyHostToDevice, st));
        const u64 *d_minc = d_in + n_queries + 1;
        bool done = false;
        if (direct) {
            // descent in the caller's order, per-query histogram + threshold writing into pinned memory: ONE wait
            u64 *host = h_in + in_words;
            host[0] = 0;
            host[1] = 0;
            if (ws->d_tmp.n == 0) ws->d_tmp.alloc(std::max<u64>(n * 16, 1 << 16));
            if (!ws->ctrl_clean || copy_form) CUDA_CHECK(cudaMemsetAsync(ws->d_ctrl.p, 0, 9 * sizeof(u64), st));
            launch_rows_kernel(m, ws, d_batch_rows, nullptr, n);
            if (hist_smem > 48 * 1024)   // (a driver call that serialises concurrent callers: only when the opt-in is needed)
                CUDA_CHECK(cudaFuncSetAttribute(query_labels_direct_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem - 1024));
            query_labels_direct_kernel<<<u32(n_queries), 256, hist_smem, st>>>(
                ws->d_row_cnt.p, ws->d_row_tmp_off.p, ws->d_tmp.p, ws->d_tmp.n, ws->d_ctrl.p, d_in, d_minc, u32(nc),
                u32(n_queries), host, cap_pairs);
            CUDA_CHECK(cudaGetLastError());
            bool flagged = false;
            if (!copy_form) {   // poll the flag the last CTA sets (bounded: a failed launch never sets it)
                volatile u64 *flag = host + 1;
                const auto t_end = std::chrono::steady_clock::now() + std::chrono::milliseconds(2);
                for (u32 spin = 0; !(flagged = *flag != 0); ++spin)
                    if ((spin & 255u) == 255u && std::chrono::steady_clock::now() > t_end) break;
                std::atomic_thread_fence(std::memory_order_acquire);
            }
            if (!flagged) CUDA_CHECK(cudaStreamSynchronize(st));
            ws->ctrl_clean = true;   // the kernel's last CTA cleared the control words
            ws->launches = 2; ws->kernel_timed = false;
            if (host[0] & 1) m->stack_safe.store(true, std::memory_order_relaxed);   // estimate too small: general path, safe mode
            bool ok = (host[0] & 3) == 0;
            u64 kept = 0;
            for (u64 q = 0; ok && q < n_queries; ++q) {
                if (host[2 + 2 * q] + host[3 + 2 * q] > cap_pairs) ok = false;
                kept += host[3 + 2 * q];
            }
            if (ok) {
                *n_out = kept;
                if (kept > capacity || (kept && !cols)) {
                    u64 o = 0;
                    for (u64 q = 0; q < n_queries; ++q) { label_offsets[q] = o; o += host[3 + 2 * q]; }
                    label_offsets[n_queries] = o;
                    rc = fail(GGQ_ERR_ARG, "capacity too small");
                } else {
                    const u64 *pairs = host + 2 + 2 * n_queries;
                    u64 o = 0;
                    for (u64 q = 0; q < n_queries; ++q) {
                        label_offsets[q] = o;
                        const u64 *src = pairs + host[2 + 2 * q];
                        const u64 k = host[3 + 2 * q];
                        for (u64 i = 0; i < k; ++i) {
                            cols[o + i] = src[i] & 0xFFFFFFFFull;
                            if (counts) counts[o + i] = src[i] >> 32;
                        }
                        o += k;
                    }
                    label_offsets[n_queries] = o;
                }
                done = true;
            } else if (host[0] & 2) {
                ws->d_tmp.alloc(std::max<u64>(n * 64, ws->d_tmp.n * 4));   // the general path grows it exactly
            }
        }
        if (!done) {
            u64 total = 0;
            if (zero_copy) {   // the general path reads the query table from device memory
                CUDA_CHECK(cudaMemcpyAsync(ws->d_qoff.p, h_in, in_words * 8, cudaMemcpyHostToDevice, st));
                d_minc = ws->d_qoff.p + n_queries + 1;
            }
            if (direct && n) CUDA_CHECK(cudaMemcpyAsync(ws->d_rows.p, d_batch_rows, n * 8, cudaMemcpyDefault, st));
            rows_device(m, ws, ws->d_rows.p, n, 4, &total);   // ordered CSR (u32 ids) in ws->d_offsets / ws->d_out
            if (ws->d_qstage_cols.n < total + 1) { ws->d_qstage_cols.alloc(total + total / 8 + 1024); ws->d_qstage_cnt.alloc(total + total / 8 + 1024); }
            if (hist_smem > 48 * 1024)
                CUDA_CHECK(cudaFuncSetAttribute(query_labels_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem - 1024));
            query_labels_kernel<<<u32(n_queries), 256, hist_smem, st>>>(ws->d_offsets.p, reinterpret_cast<const u32 *>(ws->d_out.p),
                                                                       ws->d_qoff.p, u32(nc), d_minc, ws->d_qstage_cols.p,
                                                                       ws->d_qstage_cnt.p, ws->d_qkept.p);
            CUDA_CHECK(cudaGetLastError());
            CUDA_CHECK(cudaMemsetAsync(ws->d_qkept.p + n_queries, 0, 4, st));
            size_t need = 0;
            cub::DeviceScan::ExclusiveSum(nullptr, need, ws->d_qkept.p, ws->d_qlab_off.p, n_queries + 1, st);
            if (ws->d_lab_sort.n < need + 16) ws->d_lab_sort.alloc(need + 16);
            cub::DeviceScan::ExclusiveSum(ws->d_lab_sort.p, need, ws->d_qkept.p, ws->d_qlab_off.p, n_queries + 1, st);
            CUDA_CHECK(cudaMemcpyAsync(label_offsets, ws->d_qlab_off.p, (n_queries + 1) * 8, cudaMemcpyDeviceToHost, st));
            CUDA_CHECK(cudaStreamSynchronize(st));
            const u64 kept = label_offsets[n_queries];
            *n_out = kept;
            ws->launches += 3;
            if (kept > capacity || (kept && !cols)) {
                rc = fail(GGQ_ERR_ARG, "capacity too small");
            } else if (kept) {
                if (ws->d_qout_cols.n < kept) { ws->d_qout_cols.alloc(kept + kept / 8 + 1024); ws->d_qout_cnt.alloc(kept + kept / 8 + 1024); }
                query_labels_gather_kernel<<<u32((n_queries * 8 + 255) / 256), 256, 0, st>>>(
                    ws->d_offsets.p, ws->d_qoff.p, ws->d_qlab_off.p, ws->d_qstage_cols.p, ws->d_qstage_cnt.p, n_queries,
                    ws->d_qout_cols.p, ws->d_qout_cnt.p);
                CUDA_CHECK(cudaGetLastError());
                ws->launches += 1;
                CUDA_CHECK(cudaMemcpyAsync(cols, ws->d_qout_cols.p, kept * 8, cudaMemcpyDeviceToHost, st));
                if (counts) CUDA_CHECK(cudaMemcpyAsync(counts, ws->d_qout_cnt.p, kept * 8, cudaMemcpyDeviceToHost, st));
                CUDA_CHECK(cudaStreamSynchronize(st));
            }
        }
    } catch (...) {
        m->release(ws);
        throw;
    }
    m->release(ws);
    return rc;
    GGQ_CATCH
}

int ggq_get_labels_batch_ratios(const ggq_matrix *m, const uint64_t *rows, const uint64_t *query_offsets, uint64_t n_queries,
                                const double *presence_ratios, uint64_t *label_offsets, uint64_t *cols, uint64_t *counts,
                                uint64_t capacity, uint64_t *n_out) {
    if (!query_offsets || (!presence_ratios && n_queries)) return fail(GGQ_ERR_ARG, "bad argument");
    std::vector<u64> minc;
    try { minc.resize(size_t(n_queries)); } catch (const std::bad_alloc &) { return fail(GGQ_ERR_OOM, "out of memory"); }
    for (u64 q = 0; q < n_queries; ++q) {
        if (!(presence_ratios[q] >= 0.0 && presence_ratios[q] <= 1.0)) return fail(GGQ_ERR_ARG, "presence_ratio must be in [0, 1]");
        if (query_offsets[q + 1] < query_offsets[q]) return fail(GGQ_ERR_ARG, "query_offsets must be non-decreasing");
        minc[q] = min_label_count(query_offsets[q + 1] - query_offsets[q], presence_ratios[q]);
    }
    return labels_batch_core(m, rows, query_offsets, n_queries, minc.data(), label_offsets, cols, counts, capacity, n_out);
}

int ggq_get_labels_batch(const ggq_matrix *m, const uint64_t *rows, const uint64_t *query_offsets, uint64_t n_queries,
                         double presence_ratio, uint64_t *label_offsets, uint64_t *cols, uint64_t *counts,
                         uint64_t capacity, uint64_t *n_out) {
    std::vector<double> ratios;
    try { ratios.assign(size_t(n_queries), presence_ratio); } catch (const std::bad_alloc &) { return fail(GGQ_ERR_OOM, "out of memory"); }
    return ggq_get_labels_batch_ratios(m, rows, query_offsets, n_queries, ratios.data(), label_offsets, cols, counts, capacity, n_out);
}

int ggq_count_labels_device(const ggq_matrix *m, ggq_workspace *ws, const uint64_t *d_rows, uint64_t n,
                            const uint64_t **d_counts) {
    GGQ_TRY
    if (!m || !ws || (!d_rows && n)) return fail(GGQ_ERR_ARG, "bad argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    count_labels_device(m, ws, reinterpret_cast<const u64 *>(d_rows), n);
    if (d_counts) *d_counts = reinterpret_cast<const uint64_t *>(ws->d_counts.p);
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_count_labels_allreduce_device(const ggq_matrix *m, ggq_workspace *ws, ggq_comm *comm, const uint64_t *d_rows,
                                      uint64_t n, const uint64_t **d_counts) {
    GGQ_TRY
    if (!m || !ws || !comm || (!d_rows && n)) return fail(GGQ_ERR_ARG, "bad argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    count_labels_device(m, ws, reinterpret_cast<const u64 *>(d_rows), n);
    const u64 nc = std::max<u64>(m->tree.n_cols, 1);
    // replicas: every rank counted its own slice of the batch; column shards: every rank counted its own columns
    // over the whole batch (the others are 0) — in both cases the sum over ranks is the histogram of the batch
    NCCL_CHECK(NcclApi::get().AllReduce(ws->d_counts.p, ws->d_counts.p, nc, ncclUint64, ncclSum, comm->comm, ws->stream));
    CUDA_CHECK(cudaStreamSynchronize(ws->stream));
    ws->launches += 1;
    if (d_counts) *d_counts = reinterpret_cast<const uint64_t *>(ws->d_counts.p);
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_slice_rows(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t **out, uint64_t *out_len) {
    GGQ_TRY
    if (!m || !out || !out_len || (!rows && n)) return fail(GGQ_ERR_ARG, "bad argument");
    Workspace *ws = m->acquire();
    try {
        CUDA_CHECK(cudaSetDevice(m->device));
        ws->reserve_rows(n);
        if (n) CUDA_CHECK(cudaMemcpyAsync(ws->d_rows.p, rows, n * 8, cudaMemcpyHostToDevice, ws->stream));
        u64 total = 0;
        rows_device(m, ws, ws->d_rows.p, n, 4, &total);
        u64 len = total + n;
        DeviceBuf<u64> d_slice;
        d_slice.alloc(std::max<u64>(len, 1));
        if (n) {
            slice_rows_kernel<<<u32((n + 255) / 256), 256, 0, ws->stream>>>(
                ws->d_offsets.p, reinterpret_cast<const u32 *>(ws->d_out.p), n, d_slice.p);
            CUDA_CHECK(cudaGetLastError());
        }
        std::unique_ptr<uint64_t, decltype(&std::free)> h(static_cast<uint64_t *>(std::malloc(std::max<u64>(len, 1) * 8)), &std::free);
        if (!h) throw std::bad_alloc();
        if (len) CUDA_CHECK(cudaMemcpyAsync(h.get(), d_slice.p, len * 8, cudaMemcpyDeviceToHost, ws->stream));
        CUDA_CHECK(cudaStreamSynchronize(ws->stream));
        *out = h.release();
        *out_len = len;
    } catch (...) {
        m->release(ws);
        throw;
    }
    m->release(ws);
    return GGQ_OK;
    GGQ_CATCH
}

void ggq_buffer_free(void *p) { std::free(p); }

static int get_cells(const ggq_matrix *m, const uint64_t *rows, const uint64_t *cols, u64 fixed_col, u64 n, uint8_t *out) {
    GGQ_TRY
    if (!m || (!rows && n) || (!out && n)) return fail(GGQ_ERR_ARG, "bad argument");
    if (!n) return GGQ_OK;
    if (m->kind == Matrix::BRWT && !m->tree.col_owned.empty()) {   // column-sharded: only owned columns can be answered
        const std::vector<uint8_t> &own = m->tree.col_owned;
        if (!cols && fixed_col < own.size() && !own[fixed_col]) return fail(GGQ_ERR_ARG, "column is held by another shard");
        if (cols)
            for (u64 i = 0; i < n; ++i)
                if (cols[i] < own.size() && !own[cols[i]]) return fail(GGQ_ERR_ARG, "column is held by another shard");
    }
    Workspace *ws = m->acquire();
    try {
        CUDA_CHECK(cudaSetDevice(m->device));
        ws->reserve_rows(n);
        cudaStream_t st = ws->stream;
        CUDA_CHECK(cudaMemcpyAsync(ws->d_rows.p, rows, n * 8, cudaMemcpyHostToDevice, st));
        const u64 *d_cols = nullptr;
        if (cols) {
            CUDA_CHECK(cudaMemcpyAsync(ws->d_row_tmp_off.p, cols, n * 8, cudaMemcpyHostToDevice, st));
            d_cols = ws->d_row_tmp_off.p;
        }
        unsigned char *d_res = reinterpret_cast<unsigned char *>(ws->d_bytes8.p);
        if (m->kind == Matrix::BRWT)
            if (m->wide()) brwt_get_kernel<true><<<u32((n + 255) / 256), 256, 0, st>>>(m->img, ws->d_rows.p, d_cols, fixed_col, n, d_res);
            else brwt_get_kernel<false><<<u32((n + 255) / 256), 256, 0, st>>>(m->img, ws->d_rows.p, d_cols, fixed_col, n, d_res);
        else
            colmajor_get_launch(m, ws->d_rows.p, d_cols, fixed_col, n, d_res, st);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaMemcpyAsync(out, d_res, n, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
    } catch (...) {
        m->release(ws);
        throw;
    }
    m->release(ws);
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_get(const ggq_matrix *m, const uint64_t *rows, const uint64_t *cols, uint64_t n, uint8_t *out) {
    if (!cols && n) return fail(GGQ_ERR_ARG, "null cols");
    return get_cells(m, rows, cols, 0, n, out);
}

int ggq_has_column(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t col, uint8_t *out) {
    if (m && col >= m->tree.n_cols && m->kind == Matrix::BRWT) return fail(GGQ_ERR_ARG, "column out of range");
    return get_cells(m, rows, nullptr, col, n, out);
}

// number of set rows of a column: known on the host (popcount of the column's leaf vector / of the sd vector)
static u64 column_count(const ggq_matrix *m, uint64_t col) {
    if (m->kind != Matrix::BRWT) return m->cm->col_ones[col];
    const FlatTree &t = m->tree;
    return m->lane_ones_h[t.col_path[t.col_path_off[col + 1] - 1]];
}

// BinaryMatrix::get_column on the device: ascending row ids in *d_res (enqueued on ws->stream, not awaited)
static void column_device(const ggq_matrix *m, Workspace *ws, uint64_t col, DeviceBuf<u64> *d_res, u64 *cnt_out) {
    cudaStream_t st = ws->stream;
    u64 cnt = 0;
    if (m->kind == Matrix::BRWT) {
        const FlatTree &t = m->tree;
        u32 p0 = t.col_path_off[col], p1 = t.col_path_off[col + 1];
        u32 leaf_li = t.col_path[p1 - 1];
        cnt = m->lane_ones_h[leaf_li];
        // an empty ancestor makes the column empty (BRWT.cpp:58-62); an empty leaf too
        if (d_res->n < std::max<u64>(cnt, 1)) d_res->alloc(std::max<u64>(cnt, 1));
        if (cnt) {
            u64 nb = t.group_nblocks[leaf_li >> 2];
            if (m->wide()) lane_enumerate_kernel<true><<<u32((nb + 255) / 256), 256, 0, st>>>(m->img, leaf_li, d_res->p);
            else lane_enumerate_kernel<false><<<u32((nb + 255) / 256), 256, 0, st>>>(m->img, leaf_li, d_res->p);
            for (u32 p = p1 - 1; p-- > p0;) {
                u32 li = t.col_path[p];
                if (m->lane_ones_h[li] == m->lane_size_h[li]) continue;  // all ones: identity (BRWT.cpp:77-78)
                if (m->wide()) lane_select_kernel<true><<<u32((cnt + 255) / 256), 256, 0, st>>>(m->img, li, d_res->p, cnt);
                else lane_select_kernel<false><<<u32((cnt + 255) / 256), 256, 0, st>>>(m->img, li, d_res->p, cnt);
            }
            CUDA_CHECK(cudaGetLastError());
        }
    } else {
        colmajor_get_column(m, ws, col, d_res, &cnt);
    }
    *cnt_out = cnt;
}

static int column_check(const ggq_matrix *m, uint64_t col) {
    if (col >= ggq_num_columns(m)) return fail(GGQ_ERR_ARG, "column out of range");
    if (m->kind == Matrix::BRWT && !m->tree.col_owned.empty() && !m->tree.col_owned[col])
        return fail(GGQ_ERR_ARG, "column is held by another shard");
    return GGQ_OK;
}

int ggq_get_column_count(const ggq_matrix *m, uint64_t col, uint64_t *n_rows) {
    GGQ_TRY
    if (!m || !n_rows) return fail(GGQ_ERR_ARG, "bad argument");
    if (int rc = column_check(m, col)) return rc;
    *n_rows = column_count(m, col);
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_get_column_into(const ggq_matrix *m, uint64_t col, uint64_t *rows, uint64_t capacity, uint64_t *n_rows) {
    GGQ_TRY
    if (!m || !n_rows || (!rows && capacity)) return fail(GGQ_ERR_ARG, "bad argument");
    if (int rc = column_check(m, col)) return rc;
    *n_rows = column_count(m, col);
    if (*n_rows > capacity) return fail(GGQ_ERR_ARG, "capacity too small");
    Workspace *ws = m->acquire();
    try {
        CUDA_CHECK(cudaSetDevice(m->device));
        u64 cnt = 0;
        column_device(m, ws, col, &ws->d_col, &cnt);   // workspace-owned result buffer: no cudaMalloc / cudaFree per call
        if (cnt) CUDA_CHECK(cudaMemcpyAsync(rows, ws->d_col.p, cnt * 8, cudaMemcpyDeviceToHost, ws->stream));
        CUDA_CHECK(cudaStreamSynchronize(ws->stream));
    } catch (...) {
        m->release(ws);
        throw;
    }
    m->release(ws);
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_get_column(const ggq_matrix *m, uint64_t col, uint64_t **rows, uint64_t *n_rows) {
    GGQ_TRY
    if (!m || !rows || !n_rows) return fail(GGQ_ERR_ARG, "bad argument");
    if (int rc = column_check(m, col)) return rc;
    const u64 cnt = column_count(m, col);
    uint64_t *h = static_cast<uint64_t *>(std::malloc(std::max<u64>(cnt, 1) * 8));
    if (!h) throw std::bad_alloc();
    const int rc = ggq_get_column_into(m, col, h, cnt, n_rows);
    if (rc != GGQ_OK) { std::free(h); return rc; }
    *rows = h;
    return GGQ_OK;
    GGQ_CATCH
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// ggq_get_column_noalloc: BinaryMatrix::get_column for callers that may not allocate (no counterpart in the reference).
// The K3 kernels work in a result buffer the engine normally takes from a
// workspace (cudaMalloc on first use / growth); here that scratch is a file-scope device array of a fixed, modest size
// that every GPU loading the library holds until exit.  Nothing on this path calls cudaMalloc, cudaFree or an allocating
// library routine: the kernels run on the default stream, the result leaves with a plain cudaMemcpy.
// ------------------------------------------------------------------------------------------
namespace ggq {

constexpr u64 kNoAllocMaxRows = u64(GGQ_NOALLOC_MAX_COLUMN_ROWS);   // set rows of one column: 2^20 ids = 8 MiB of scratch
static __device__ __align__(256) unsigned long long g_noalloc_column[kNoAllocMaxRows];

// In host code `g_noalloc_column` is only a handle of the symbol, not a device address: the address is looked up once
// per device, with that device current, and the POINTER is what the kernels get.
constexpr int kNoAllocMaxDevices = 64;
static std::mutex g_noalloc_mu;              // also serialises the calls of one process: they all share the array
static u64 *g_noalloc_ptr[kNoAllocMaxDevices];

static u64 *noalloc_scratch(int device) {    // caller holds g_noalloc_mu and has made `device` current
    if (device < 0 || device >= kNoAllocMaxDevices) throw LimitError("device ordinal beyond the no-alloc table");
    if (!g_noalloc_ptr[device]) {
        void *p = nullptr;
        CUDA_CHECK(cudaGetSymbolAddress(&p, g_noalloc_column));
        if (!p) throw CudaError("cudaGetSymbolAddress returned no address for the no-alloc scratch");
        g_noalloc_ptr[device] = static_cast<u64 *>(p);
    }
    return g_noalloc_ptr[device];
}

}  // namespace ggq

extern "C" int ggq_get_column_noalloc(const ggq_matrix *m, uint64_t col, uint64_t *rows, uint64_t capacity, uint64_t *n_rows) {
    GGQ_TRY
    if (!m || !n_rows || (!rows && capacity)) return fail(GGQ_ERR_ARG, "bad argument");
    if (m->kind != Matrix::BRWT) return fail(GGQ_ERR_ARG, "ggq_get_column_noalloc: BRWT matrices only");
    if (int rc = column_check(m, col)) return rc;
    const u64 cnt = column_count(m, col);
    *n_rows = cnt;
    if (cnt > kNoAllocMaxRows)
        return fail(GGQ_ERR_LIMIT, "ggq_get_column_noalloc: the column has " + std::to_string(cnt) + " set rows, the fixed scratch holds " +
                                   std::to_string(kNoAllocMaxRows) + " (GGQ_NOALLOC_MAX_COLUMN_ROWS): use ggq_get_column_into");
    if (cnt > capacity) return fail(GGQ_ERR_ARG, "capacity too small");
    if (!cnt) return GGQ_OK;   // an empty leaf or ancestor (BRWT.cpp:58-62)
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    CUDA_CHECK(cudaSetDevice(m->device));
    u64 *d_res = noalloc_scratch(m->device);
    // Entries [0, cnt) are all written by the enumeration before anything reads them: what the previous call left in the
    // array needs no clearing.
    const FlatTree &t = m->tree;
    const u32 p0 = t.col_path_off[col], p1 = t.col_path_off[col + 1];
    const u32 leaf_li = t.col_path[p1 - 1];
    const u64 nb = t.group_nblocks[leaf_li >> 2];
    if (m->wide()) lane_enumerate_kernel<true><<<u32((nb + 255) / 256), 256>>>(m->img, leaf_li, d_res);
    else lane_enumerate_kernel<false><<<u32((nb + 255) / 256), 256>>>(m->img, leaf_li, d_res);
    for (u32 p = p1 - 1; p-- > p0;) {
        const u32 li = t.col_path[p];
        if (m->lane_ones_h[li] == m->lane_size_h[li]) continue;   // all ones: identity (BRWT.cpp:77-78)
        if (m->wide()) lane_select_kernel<true><<<u32((cnt + 255) / 256), 256>>>(m->img, li, d_res, cnt);
        else lane_select_kernel<false><<<u32((cnt + 255) / 256), 256>>>(m->img, li, d_res, cnt);
    }
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaMemcpy(rows, d_res, cnt * 8, cudaMemcpyDeviceToHost));   // (default stream: ordered after the kernels)
    return GGQ_OK;
    GGQ_CATCH
}

#include "ggq_bv_api.cuh"
