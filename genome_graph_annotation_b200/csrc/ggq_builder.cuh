// K6 (benchmark harness): build, ON THE DEVICE, the BRWT that
//   BRWTBottomUpBuilder::build(columns, get_basic_partitioner(arity))   (BRWT_builders.cpp:19-31,122-163)
// produces for the implicit matrix M[i][j] = mix64(seed*0x9E3779B97F4A7C15 + i*n_cols + j) < threshold
// (SURVEY.md §8d), directly in the slot layout.  The reference builder needs the materialised columns
// (compute_or :33-49, generate_subindex :51-66); at 3.7e9 x 2652 that is 1.2 TB, so the matrix is never
// materialised here: a warp owns 32 consecutive rows (one per lane) and interprets a post-order
// "program" of the tree shape, keeping one bit per open node in a 64-bit register stack:
//     LEAF c   : push M[row][c]
//     NODE a   : the top a entries are the children's OR bits; P = OR of them; every child's bit is
//                appended to the child's vector at the rows where P is set (the compaction of
//                generate_subindex, done with __ballot_sync/__popc), then the a entries are replaced by P.
// Pass 1 only counts P per (tile of 4096 rows, node); an exclusive scan over tiles gives every tile its
// write cursor in every vector; pass 2 re-evaluates the hashes and ORs the compacted bits in place.
#pragma once
#include "ggq_engine.hpp"
#include "ggq_kernels.cuh"

namespace ggq {

constexpr u32 kTileRows = 4096;
constexpr int kBuildWarps = 4;

struct BuildInstr {
    u32 op_arity;   // bit 0: 1 = NODE; bits 8..: arity (NODE)
    u32 arg;        // LEAF: column; NODE: index of the first child in child_li[]
    u32 node_idx;   // NODE: ordinal among internal nodes; LEAF: ordinal among leaves
    u32 pad;
};

struct BuildArgs {
    const BuildInstr *prog;
    u32 n_instr;
    const u32 *child_li;     // lane (group*4+lane) of every child, in program order
    u32 n_internal;
    u32 n_counters;          // counters per tile: n_internal, or n_internal + leaves when leaves are counted too
    u64 n_rows, n_cols, threshold, seed_mul;
    u64 n_tiles;
    unsigned short *tile_cnt;  // [tile][node_idx]   (pass 1 out)
    const u64 *tile_prefix;    // [tile][node_idx]   (pass 2 in)
    uint4 *slots;
    const u64 *group_hot;
    u32 dw, quads;             // data words and 16-byte units per slot (3 / 1, or 7 / 2 in the capacity layout)
};

template <bool FILL>
__global__ void __launch_bounds__(kBuildWarps * 32) brwt_build_kernel(BuildArgs a) {
    extern __shared__ u32 s_cnt_all[];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const u32 lt_mask = (1u << lane) - 1u;
    u32 *s_cnt = s_cnt_all + size_t(warp) * a.n_counters;
    for (u64 tile = u64(blockIdx.x) * kBuildWarps + warp; tile < a.n_tiles; tile += u64(gridDim.x) * kBuildWarps) {
        for (u32 i = lane; i < a.n_counters; i += 32) s_cnt[i] = 0;
        __syncwarp();
        const u64 *prefix = FILL ? a.tile_prefix + tile * a.n_internal : nullptr;
        for (u32 step = 0; step < kTileRows / 32; ++step) {
            const u64 row = tile * kTileRows + step * 32 + lane;
            const bool row_ok = row < a.n_rows;
            if (!__any_sync(0xffffffffu, row_ok)) break;
            const u64 hbase = a.seed_mul + row * a.n_cols;
            u64 stack = 0;
            u32 sp = 0;
            for (u32 pc = 0; pc < a.n_instr; ++pc) {
                const BuildInstr ins = a.prog[pc];
                if (!(ins.op_arity & 1u)) {
                    u64 bit = (row_ok && mix64(hbase + ins.arg) < a.threshold) ? 1ull : 0ull;
                    stack |= bit << sp;
                    ++sp;
                    if (!FILL && a.n_counters > a.n_internal) {   // count pass of the shape search: column popcounts too
                        const u32 c = __popc(__ballot_sync(0xffffffffu, bit != 0));
                        if (lane == 0 && c) s_cnt[a.n_internal + ins.node_idx] += c;
                        __syncwarp();
                    }
                } else {
                    const u32 ar = ins.op_arity >> 8;
                    sp -= ar;
                    const u32 kids = u32(stack >> sp) & ((1u << ar) - 1u);
                    const bool P = kids != 0;
                    stack &= (1ull << sp) - 1ull;
                    stack |= u64(P) << sp;
                    ++sp;
                    const u32 pm = __ballot_sync(0xffffffffu, P);
                    if (pm) {
                        const u32 before = s_cnt[ins.node_idx];
                        if (FILL) {
                            const u64 g0 = prefix[ins.node_idx] + before;   // first target position in every child
                            const u64 gp = g0 + __popc(pm & lt_mask);
                            const u64 wa = g0 >> 5;
                            const bool in_a = P && (gp >> 5) == wa;
                            const bool in_b = P && !in_a;
                            const u32 sh = u32(gp) & 31u;
                            const bool two = ((g0 + __popc(pm) - 1) >> 5) != wa;
                            for (u32 k = 0; k < ar; ++k) {
                                const u32 cb = (kids >> k) & 1u;
                                u32 va = __reduce_or_sync(0xffffffffu, in_a ? (cb << sh) : 0u);
                                u32 vb = two ? __reduce_or_sync(0xffffffffu, in_b ? (cb << sh) : 0u) : 0u;
                                if (lane == 0 && (va | vb)) {
                                    const u32 li = a.child_li[ins.arg + k];
                                    if (li == ~0u) continue;
                                    const u64 hot = a.group_hot[li >> 2];
                                    const u32 sl = u32(hot >> kGhStrideShift) & 3u;
                                    uint4 *base = a.slots + (hot & kGhBaseMask) + (li & 3) * a.quads;
                                    if (va) {
                                        u64 blk = a.dw == 3 ? wa / 3 : wa / 7; u32 w = u32(wa - blk * a.dw);
                                        atomicOr(reinterpret_cast<u32 *>(base + (blk << sl)) + 1 + w, va);
                                    }
                                    if (vb) {
                                        u64 wb = wa + 1;
                                        u64 blk = a.dw == 3 ? wb / 3 : wb / 7; u32 w = u32(wb - blk * a.dw);
                                        atomicOr(reinterpret_cast<u32 *>(base + (blk << sl)) + 1 + w, vb);
                                    }
                                }
                            }
                        }
                        __syncwarp();
                        if (lane == 0) s_cnt[ins.node_idx] = before + __popc(pm);
                        __syncwarp();
                    }
                }
            }
            if (FILL) {
                // the root's own vector: position = row, one aligned word per 32 rows
                const u32 rootbits = __ballot_sync(0xffffffffu, (stack & 1ull) != 0);
                if (lane == 0 && rootbits) {
                    const u64 hot = a.group_hot[0];
                    const u32 sl = u32(hot >> kGhStrideShift) & 3u;
                    const u64 wr = (tile * kTileRows + step * 32) >> 5;
                    const u64 blk = a.dw == 3 ? wr / 3 : wr / 7; const u32 w = u32(wr - blk * a.dw);
                    reinterpret_cast<u32 *>(a.slots + (hot & kGhBaseMask) + (blk << sl))[1 + w] = rootbits;
                }
            }
        }
        if (!FILL) {
            __syncwarp();
            for (u32 i = lane; i < a.n_counters; i += 32)
                a.tile_cnt[tile * a.n_counters + i] = (unsigned short)s_cnt[i];
        }
        __syncwarp();
    }
}

// exclusive scan over tiles, one thread per internal node (coalesced across nodes)
__global__ void tile_scan_kernel(const unsigned short *tile_cnt, u64 *tile_prefix, u64 *node_ones,
                                 u64 n_tiles, u32 n_internal) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_internal) return;
    u64 run = 0;
    for (u64 t = 0; t < n_tiles; ++t) {
        u64 c = tile_cnt[t * n_internal + i];
        tile_prefix[t * n_internal + i] = run;
        run += c;
    }
    node_ones[i] = run;
}

// totals per counter over all tiles (count-only pass)
__global__ void tile_sum_kernel(const unsigned short *tile_cnt, u64 *totals, u64 n_tiles, u32 n_counters) {
    u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_counters) return;
    u64 run = 0;
    for (u64 t = 0; t < n_tiles; ++t) run += tile_cnt[t * n_counters + i];
    totals[i] = run;
}

// `shape`: FlatTree::nodes (parent / children / cols / depth) of the tree to build.  count_only != null: only evaluate
// the popcount of EVERY node's index vector (internal nodes and leaves) into (*count_only)[node id] and return null —
// what BRWTOptimizer::relax needs before the final shape is known.
inline Matrix *synthesize_brwt_shape(FlatTree &&shape, u64 n_rows, u32 n_cols, u64 threshold, u64 seed, int device,
                                     u32 shard = 0, u32 n_shards = 1, std::vector<u64> *count_only = nullptr) {
    std::unique_ptr<Matrix> m(new Matrix());
    m->kind = Matrix::BRWT;
    m->device = device;
    CUDA_CHECK(cudaSetDevice(device));
    m->tree = std::move(shape);
    FlatTree &t = m->tree;
    t.slot_bits = default_slot_bits();
    assign_groups(&t);
    plan_shards(&t, shard, n_shards);
    if (n_rows >= kMaxRows) throw LimitError("too many rows");
    const bool root_is_leaf = t.nodes[0].children.empty();

    // post-order program
    std::vector<BuildInstr> prog;
    std::vector<u32> child_li;
    std::vector<int> internal_of(t.nodes.size(), -1);
    std::vector<u32> internal_nodes, leaf_nodes;
    u32 max_sp = 0;
    if (n_cols) {
        struct Frame { int id; size_t next; };
        std::vector<Frame> st;
        st.push_back({0, 0});
        u32 sp = 0;
        while (!st.empty()) {
            Frame &f = st.back();
            const FlatNode &n = t.nodes[f.id];
            if (n.children.empty()) {
                prog.push_back({0u, n.cols[0], u32(leaf_nodes.size()), 0u});
                leaf_nodes.push_back(u32(f.id));
                max_sp = std::max(max_sp, ++sp);
                st.pop_back();
            } else if (f.next < n.children.size()) {
                int c = n.children[f.next++];
                st.push_back({c, 0});
            } else {
                u32 ar = u32(n.children.size());
                if (ar > 31) throw LimitError("arity too large for the device builder");
                internal_of[f.id] = int(internal_nodes.size());
                prog.push_back({1u | (ar << 8), u32(child_li.size()), u32(internal_nodes.size()), 0u});
                internal_nodes.push_back(u32(f.id));
                for (int c : n.children)   // ~0u: vector held by another shard — evaluated for the OR, never written
                    child_li.push_back(t.present(u32(c)) ? t.nodes[c].group * 4 + t.nodes[c].lane : ~0u);
                sp -= ar - 1;
                st.pop_back();
            }
        }
    }
    if (max_sp > 64) throw LimitError("tree too deep for the device builder");
    const u32 NI = u32(internal_nodes.size());
    const u32 NC = count_only ? NI + u32(leaf_nodes.size()) : NI;   // counters per tile
    const u64 n_tiles = (n_rows + kTileRows - 1) / kTileRows;

    DeviceBuf<BuildInstr> d_prog;
    DeviceBuf<u32> d_child;
    DeviceBuf<unsigned short> d_cnt;
    DeviceBuf<u64> d_prefix, d_ones;
    d_prog.alloc(std::max<size_t>(prog.size(), 1));
    if (!prog.empty()) CUDA_CHECK(cudaMemcpy(d_prog.p, prog.data(), prog.size() * sizeof(BuildInstr), cudaMemcpyHostToDevice));
    d_child.alloc(std::max<size_t>(child_li.size(), 1));
    if (!child_li.empty()) CUDA_CHECK(cudaMemcpy(d_child.p, child_li.data(), child_li.size() * 4, cudaMemcpyHostToDevice));
    d_cnt.alloc(std::max<u64>(n_tiles * NC, 1));
    if (!count_only) d_prefix.alloc(std::max<u64>(n_tiles * NI, 1));
    d_ones.alloc(std::max<u32>(NC, 1));

    BuildArgs a{};
    a.prog = d_prog.p; a.n_instr = u32(prog.size()); a.child_li = d_child.p; a.n_internal = NI; a.n_counters = NC;
    a.n_rows = n_rows; a.n_cols = n_cols; a.threshold = threshold;
    a.seed_mul = seed * 0x9E3779B97F4A7C15ULL;
    a.n_tiles = n_tiles; a.tile_cnt = d_cnt.p; a.tile_prefix = d_prefix.p;
    int sms = 148;
    CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    size_t smem = size_t(kBuildWarps) * std::max<u32>(NC, 1) * 4;
    {
        int max_smem = 0;
        CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
        if (smem > size_t(max_smem)) throw LimitError("too many internal nodes for the device builder (benchmark harness)");
    }
    CUDA_CHECK(cudaFuncSetAttribute(brwt_build_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    CUDA_CHECK(cudaFuncSetAttribute(brwt_build_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    int per_sm = 1;
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, brwt_build_kernel<true>, kBuildWarps * 32, smem));
    u32 grid = u32(std::min<u64>(u64(sms) * std::max(per_sm, 1), (n_tiles + kBuildWarps - 1) / kBuildWarps));
    if (grid == 0) grid = 1;

    if (count_only) {
        count_only->assign(t.nodes.size(), 0);
        if (n_tiles && NC) {
            brwt_build_kernel<false><<<grid, kBuildWarps * 32, smem>>>(a);
            CUDA_CHECK(cudaGetLastError());
            tile_sum_kernel<<<(NC + 127) / 128, 128>>>(d_cnt.p, d_ones.p, n_tiles, NC);
            CUDA_CHECK(cudaGetLastError());
            std::vector<u64> tot(NC);
            CUDA_CHECK(cudaMemcpy(tot.data(), d_ones.p, size_t(NC) * 8, cudaMemcpyDeviceToHost));
            for (u32 k = 0; k < NI; ++k) (*count_only)[internal_nodes[k]] = tot[k];
            for (size_t k = 0; k < leaf_nodes.size(); ++k) (*count_only)[leaf_nodes[k]] = tot[NI + k];
        }
        return nullptr;
    }
    std::vector<u64> ones(std::max<u32>(NI, 1), 0);
    if (n_tiles && NI) {
        brwt_build_kernel<false><<<grid, kBuildWarps * 32, smem>>>(a);
        CUDA_CHECK(cudaGetLastError());
        tile_scan_kernel<<<(NI + 127) / 128, 128>>>(d_cnt.p, d_prefix.p, d_ones.p, n_tiles, NI);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaMemcpy(ones.data(), d_ones.p, NI * 8, cudaMemcpyDeviceToHost));
    }
    d_cnt.release();
    // sizes: a child is as long as its parent has ones
    for (u32 k = 0; k < NI; ++k) t.nodes[internal_nodes[k]].ones = ones[k];
    for (size_t id = 1; id < t.nodes.size(); ++id) t.nodes[id].size = t.nodes[t.nodes[id].parent].ones;
    assign_geometry(&t);
    m->upload_tables(t);
    m->slots.alloc(std::max<u64>(t.total_slots, 4));
    CUDA_CHECK(cudaMemset(m->slots.p, 0, m->slots.bytes()));
    a.slots = m->slots.p;
    a.group_hot = m->group_hot.p;
    a.dw = slot_data_words(t.slot_bits);
    a.quads = slot_quads(t.slot_bits);
    if (n_tiles && n_cols) {
        if (root_is_leaf) {
            // single column: the program is one LEAF; pass 2 only writes the root vector
        }
        brwt_build_kernel<true><<<grid, kBuildWarps * 32, smem>>>(a);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaDeviceSynchronize());
    }
    d_prefix.release();
    m->sel.alloc(1);
    m->finalize_arena(t, false);
    m->refresh_img();
    m->build_leaf_codes();
    m->derive_stats();
    return m.release();
}

inline Matrix *synthesize_brwt(u64 n_rows, u32 n_cols, u64 threshold, u64 seed, u32 arity, int device,
                               u32 shard = 0, u32 n_shards = 1) {
    FlatTree t;
    build_basic_shape(n_rows, n_cols, arity, &t);
    return synthesize_brwt_shape(std::move(t), n_rows, n_cols, threshold, seed, device, shard, n_shards);
}

}  // namespace ggq
