// Parity-grade BRWT builder on the device: the reference's bottom-up construction
//   BRWTBottomUpBuilder::build(columns, partitioner)              annotation/hierarchical_annotation/BRWT_builders.cpp:122-163
//   partitioner = parallel_binary_grouping_greedy                  .../partitionings.cpp:146-196
//               | get_basic_partitioner(arity)                     .../BRWT_builders.cpp:19-31
//   BRWTOptimizer::relax(max_arity)                                .../BRWT_builders.cpp:166-380
// with the heavy parts on the GPU and the decisions taken exactly as the reference takes them, so that the SAME tree
// comes out (same nodes, same child order, same column arrangement — checked node by node against the reference
// builder in tests/test_builder_gpu.py):
//
//   * similarity of two columns = popcount(a & b) over (up to) 10^6 sampled rows (partitionings.cpp:73-96,128-143).
//     The sample is drawn with the reference's sequence of std:: random calls (utils.cpp:777-816, seed 1), on the host;
//     the Gram matrix — a bit-matrix product, the one contraction on this path — is a tiled AND+POPC kernel;
//   * pairs are ranked with std::sort under the reference's comparator (similarity, then distance in the current
//     arrangement) on the same candidate order, so ties fall where libstdc++ lets them fall in the reference too;
//   * a parent's index vector is the OR of its children (compute_or, BRWT_builders.cpp:33-49): one kernel;
//   * relax needs only (size, popcount) of every node vector and the reference's RRR space model
//     (bv_space_taken_rrr / pruning_delta, BRWT_builders.cpp:281-380), evaluated in double in the reference's order;
//   * a stored child vector is the child's index vector restricted to the ones of its parent
//     (generate_subindex, BRWT_builders.cpp:51-66): a parallel bit-extract (pext) kernel writing the plain bits the
//     slot packer consumes.
#pragma once
#include <cmath>
#include <functional>
#include <random>

namespace ggq {

// ---- kernels ---------------------------------------------------------------------------------------------------

// G[j * m + k] = popcount(v_j & v_k) over W words, for k < j (other entries untouched).  One CTA computes a 64 x 64
// tile of pairs; the 128 column chunks of 32 words each are staged in shared memory (32 KB), every thread keeps a
// 4 x 4 block of partial counts in registers.
constexpr int kGramTile = 64, kGramChunk = 32;
__global__ void __launch_bounds__(256)
gram_kernel(const u64 *const *vecs, u32 m, u64 W, u32 *G) {
    __shared__ u64 sj[kGramTile][kGramChunk + 1], sk[kGramTile][kGramChunk + 1];
    const u32 tj = blockIdx.y, tk = blockIdx.x;
    if (tk > tj) return;   // only the lower triangle
    const u32 j0 = tj * kGramTile, k0 = tk * kGramTile;
    const u32 tx = threadIdx.x & 15, ty = threadIdx.x >> 4;   // 16 x 16 threads, 4 x 4 pairs each
    u32 acc[4][4] = {};
    for (u64 w0 = 0; w0 < W; w0 += kGramChunk) {
        for (u32 i = threadIdx.x; i < kGramTile * kGramChunk; i += 256) {
            const u32 c = i / kGramChunk, w = i % kGramChunk;
            const u64 wi = w0 + w;
            sj[c][w] = (j0 + c < m && wi < W) ? vecs[j0 + c][wi] : 0ull;
            sk[c][w] = (k0 + c < m && wi < W) ? vecs[k0 + c][wi] : 0ull;
        }
        __syncthreads();
        #pragma unroll 4
        for (u32 w = 0; w < kGramChunk; ++w) {
            u64 a[4], b[4];
            #pragma unroll
            for (u32 i = 0; i < 4; ++i) { a[i] = sj[ty * 4 + i][w]; b[i] = sk[tx * 4 + i][w]; }
            #pragma unroll
            for (u32 i = 0; i < 4; ++i)
                #pragma unroll
                for (u32 l = 0; l < 4; ++l) acc[i][l] += __popcll(a[i] & b[l]);
        }
        __syncthreads();
    }
    for (u32 i = 0; i < 4; ++i)
        for (u32 l = 0; l < 4; ++l) {
            const u32 j = j0 + ty * 4 + i, k = k0 + tx * 4 + l;
            if (j < m && k < j) G[size_t(j) * m + k] = acc[i][l];
        }
}

__global__ void or_into_kernel(u64 *dst, const u64 *src, u64 W) {
    const u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < W) dst[i] |= src[i];
}

// sub bit i = full bit idx[i]
__global__ void sample_rows_kernel(const u64 *full, const u64 *idx, u64 n_idx, u64 *sub) {
    const u64 w = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (w * 64 >= n_idx) return;
    u64 x = 0;
    for (u32 b = 0; b < 64 && w * 64 + b < n_idx; ++b) {
        const u64 r = idx[w * 64 + b];
        x |= ((full[r >> 6] >> (r & 63)) & 1ull) << b;
    }
    sub[w] = x;
}

__global__ void word_popc_kernel(const u64 *v, u64 W, u64 *cnt) {
    const u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < W) cnt[i] = u64(__popcll(v[i]));
}

// out (plain bits, u32 words, zeroed) = child restricted to the ones of parent; before[w] = ones of parent before word w
__global__ void pext_compact_kernel(const u64 *child, const u64 *parent, const u64 *before, u64 W, u32 *out) {
    const u64 w = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (w >= W) return;
    u64 mask = parent[w];
    const u64 c = child[w];
    if (!(c & mask)) return;
    u64 bits = 0;
    u32 k = 0;
    while (mask) {   // software pext
        const u64 low = mask & (~mask + 1);
        if (c & low) bits |= 1ull << k;
        ++k;
        mask &= mask - 1;
    }
    const u64 pos = before[w];
    const u64 wi = pos >> 5;
    const u32 off = u32(pos & 31);
    atomicOr(out + wi, u32(bits << off));
    const u64 rest = off ? bits >> (32 - off) : bits >> 32;   // (off == 0: the low word took 32 bits)
    if (off) {
        if (u32(rest)) atomicOr(out + wi + 1, u32(rest));
        if (rest >> 32) atomicOr(out + wi + 2, u32(rest >> 32));
    } else if (u32(rest)) {
        atomicOr(out + wi + 1, u32(rest));
    }
}

// plain u64 words -> plain u32 words (the root vector is stored as it is)
__global__ void words64_to_32_kernel(const u64 *in, u64 W, u32 *out) {
    const u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < W) { out[2 * i] = u32(in[i]); out[2 * i + 1] = u32(in[i] >> 32); }
}

// ---- host side ---------------------------------------------------------------------------------------------------

constexpr u64 kBuildRowsSampled = 1000000;   // kNumRowsSampled, partitionings.cpp:5

// Row sample of estimate_similarities -> random_submatrix(columns, min(10^6, n), seed 1) (partitionings.cpp:35-56,128-143):
// utils::sample_indexes (utils.cpp:777-816) consumes a std::mt19937 in a fixed sequence of std:: distribution calls;
// the same calls on the same libstdc++ give the same rows.  Returned sorted, as random_submatrix sorts them.
inline std::vector<u64> build_row_sample(u64 n_rows) {
    if (!n_rows) return {};
    const u64 want = std::min<u64>(n_rows, kBuildRowsSampled);
    std::mt19937 gen;
    gen.seed(1);
    std::vector<u64> idx;
    idx.reserve(size_t(3 * want));
    if (want * 10 < n_rows) {   // sparse sample: draw 1.5 x as many, dedupe, repeat until enough
        std::uniform_int_distribution<u64> pick(0, n_rows - 1);
        while (idx.size() < want) {
            idx.clear();
            for (size_t i = 0; i < 1.5 * want; ++i) idx.push_back(pick(gen));
            std::sort(idx.begin(), idx.end());
            idx.erase(std::unique(idx.begin(), idx.end()), idx.end());
        }
    } else {                    // dense sample: every row with probability 2 * want / n
        std::bernoulli_distribution keep(2.0 * want / n_rows);
        while (idx.size() < want) {
            idx.clear();
            for (u64 i = 0; i < n_rows; ++i)
                if (keep(gen)) idx.push_back(i);
        }
    }
    std::shuffle(idx.begin(), idx.end(), gen);
    idx.resize(size_t(want));
    std::sort(idx.begin(), idx.end());
    return idx;
}

struct BuildNode {
    std::vector<int> children;   // node ids, stored (= DFS) order
    std::vector<u32> cols;       // global columns in the node's arrangement (children's arrangements concatenated)
    u64 ones = 0;                // popcount of the node's index vector (OR of its columns)
};

// Greedy pairing of the current level (parallel_binary_grouping_greedy, partitionings.cpp:146-196) from its Gram
// matrix G (row-major, G[j * k_n + k] valid for k < j).  Groups are {larger index, smaller index} in candidate order,
// then the unmatched positions in increasing order.
inline std::vector<std::vector<u32>> greedy_partition(const std::vector<u32> &G, u32 k_n) {
    struct Cand { size_t a, b; u64 sim; };
    std::vector<Cand> cand;
    cand.reserve(size_t(k_n) * (k_n - 1) / 2);
    for (size_t j = 1; j < k_n; ++j)
        for (size_t k = 0; k < j; ++k) cand.push_back({j, k, u64(G[j * k_n + k])});
    auto gap = [](size_t x, size_t y) { return x > y ? x - y : y - x; };
    // std::sort, not stable_sort: the reference leaves full ties to the sort's own order, and so must this
    std::sort(cand.begin(), cand.end(), [&](const Cand &p, const Cand &q) {
        return p.sim > q.sim || (p.sim == q.sim && gap(p.a, p.b) < gap(q.a, q.b));
    });
    std::vector<std::vector<u32>> part;
    part.reserve((k_n + 1) / 2);
    std::vector<bool> used(k_n, false);
    for (const Cand &c : cand)
        if (!used[c.a] && !used[c.b]) {
            used[c.a] = used[c.b] = true;
            part.push_back({u32(c.a), u32(c.b)});
        }
    for (u32 i = 0; i < k_n; ++i)
        if (!used[i]) part.push_back({i});
    return part;
}

inline std::vector<std::vector<u32>> basic_partition(u32 k_n, u32 arity) {
    std::vector<std::vector<u32>> part((k_n + arity - 1) / arity);
    for (u32 i = 0; i < k_n; ++i) part[i / arity].push_back(i);
    return part;
}

// RRR space model of BRWTOptimizer (BRWT_builders.cpp:281-300): bits taken by an rrr_vector<63> of `size` bits with
// `ones` set.  216 = sizeof(bit_vector_rrr<>) in the reference build (checked by tests/test_oracle_cpu.py).
constexpr double kSizeofBitVectorRrrBits = 216 * 8;
inline double rrr_space_bits(u64 size, u64 ones) {
    const double logbinomial = (std::lgamma(size + 1) - std::lgamma(ones + 1) - std::lgamma(size - ones + 1)) / std::log(2);
    return logbinomial + std::ceil(std::log2(63 + 1) / 63) * size + kSizeofBitVectorRrrBits;
}

// BRWTOptimizer::relax (BRWT_builders.cpp:166-211) on the node statistics: internal nodes in reverse BFS order; a
// child is dissolved into its parent when it is internal, the arity allows it and pruning_delta (BRWT_builders.cpp:
// 352-380) is not positive.  size_of[x] = length of x's stored vector = ones of its current parent (root: n_rows).
inline void relax_tree(std::vector<BuildNode> &N, int root, u64 n_rows, u64 max_arity) {
    std::vector<int> internal;   // BFS order
    {
        std::vector<int> queue = {root};
        for (size_t h = 0; h < queue.size(); ++h) {
            const int id = queue[h];
            if (!N[id].children.empty()) internal.push_back(id);
            for (int c : N[id].children) queue.push_back(c);
        }
    }
    std::vector<u64> size_of(N.size(), 0);
    size_of[root] = n_rows;
    {
        std::vector<int> queue = {root};
        for (size_t h = 0; h < queue.size(); ++h)
            for (int c : N[queue[h]].children) { size_of[c] = N[queue[h]].ones; queue.push_back(c); }
    }
    auto pruning_delta = [&](int node) {
        double delta = 0;
        for (int x : N[node].children) {
            delta += rrr_space_bits(size_of[node], N[x].ones);   // the child's vector, re-indexed by the grandparent
            delta -= rrr_space_bits(size_of[x], N[x].ones);      // its old vector
        }
        delta -= rrr_space_bits(size_of[node], N[node].ones);    // the removed index vector
        return delta;
    };
    for (size_t h = internal.size(); h-- > 0;) {   // leaves first, root last
        const int p = internal[h];
        const std::vector<int> old = N[p].children;
        std::vector<int> now;
        for (size_t g = 0; g < old.size(); ++g) {
            const int c = old[g];
            const u64 room = max_arity - std::min<u64>(max_arity, u64(now.size()) + old.size() - g - 1);
            bool prune = !N[c].children.empty() && N[c].children.size() <= room;
            if (prune && pruning_delta(c) > 0) prune = false;
            if (!prune) { now.push_back(c); continue; }
            for (int x : N[c].children) { size_of[x] = size_of[c]; now.push_back(x); }   // reassign (BRWT_builders.cpp:239-279)
            N[c].children.clear();
        }
        N[p].children = now;
    }
}

}  // namespace ggq
