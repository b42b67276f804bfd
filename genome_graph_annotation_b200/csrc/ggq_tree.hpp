// Host-side flattening of a BRWT into the level-order, sibling-interleaved device image plan.
//
// Input: the pre-order node stream of BRWT::load (BRWT.cpp:87-111) parsed by ggq_format.hpp, or a
// synthetic tree shape (basic partitioner, BRWT_builders.cpp:19-31 + build loop :122-163).
// Output (FlatTree): nodes renumbered breadth-first (the order BRWT::BFT visits them, BRWT.cpp:204-220),
// the global column of every leaf obtained by composing assignments_.get() up to the root
// (BRWT.cpp:49), the DFS (= get_row) rank of every leaf column, the root-to-leaf path of every
// column (assignments_.group()/rank(), BRWT.cpp:21-23,73-74), and the slot arena geometry.
#pragma once
#include <atomic>
#include <thread>

#include <algorithm>
#include <cstdint>
#include <deque>
#include <vector>

#include "ggq_format.hpp"
#include "ggq_layout.hpp"

namespace ggq {

struct FlatNode {
    int parent = -1;
    int child_index = 0;             // position among the parent's children
    int depth = 0;
    std::vector<int> children;       // BFS ids, stored (= DFS) order
    uint64_t size = 0;               // bits of nonzero_rows_
    uint64_t ones = 0;               // popcount (filled by host parse; by the device for synthetic trees)
    std::vector<uint32_t> cols;      // global column id of every local column
    uint32_t group = 0, lane = 0;    // where the vector lives
    uint32_t first_child_group = 0, n_child_groups = 0;
    uint32_t leaf_rank = 0;          // leaves: DFS rank of the first column
    PlainBits bits;                  // empty for synthetic trees
    RrrRaw raw;                      // lazy load: the undecoded vector (bits.w empty until upload decodes it)
    bool lazy = false;
    std::vector<std::vector<uint32_t>> partition;   // parsed trees: the node's RangePartition as it was in the stream
};

struct FlatTree {
    uint64_t n_rows = 0, n_cols = 0;
    std::vector<FlatNode> nodes;
    // geometry
    uint32_t slot_bits = kSlotBits;       // kSlotBits (default) or kWideSlotBits (capacity layout), set before assign_groups
    std::vector<uint64_t> group_hot;      // see ggq_layout.hpp
    std::vector<uint64_t> group_nblocks;  // incl. sentinel
    std::vector<uint32_t> group_parent;   // node whose children the group holds (root group: ~0u)
    std::vector<uint32_t> lane_info, lane_aux, lane_node;  // size 4 * n_groups; lane_node = ~0u if unused
    std::vector<uint32_t> lane_sb_off;    // first superblock base of the lane
    std::vector<uint64_t> lane_sel_off;   // first select sample of the lane
    std::vector<uint32_t> leaf_col;       // DFS rank -> global column
    std::vector<uint32_t> col_path_off;   // n_cols + 1
    std::vector<uint32_t> col_path;       // (group << 2 | lane), root first
    // leaf codes (ggq_rows.cuh): per lane, 1 if the lane's node carries codes; code_rec[li] = {offset lo, hi, rank0, leaves}
    std::vector<uint8_t> lane_coded;
    std::vector<uint32_t> code_rec;       // 4 words per lane (empty when no lane is coded)
    uint64_t total_slots = 0, total_sb = 0, total_sel = 0;
    uint32_t max_depth = 0;
    bool multi_column_leaves = false;
    // column sharding: nodes above the cut are replicated, subtrees below it are dealt to shards
    uint32_t shard = 0, n_shards = 1, cut_depth = 0;
    std::vector<uint8_t> node_present;    // empty = everything present
    std::vector<uint8_t> col_owned;       // per global column; empty = all
    std::vector<uint32_t> col_owner;      // per global column: owning shard (filled by plan_shards)
    bool present(uint32_t node) const { return node_present.empty() || node_present[node]; }

    uint32_t n_groups() const { return uint32_t(group_hot.size()); }
};

// Decodes the lazily parsed vectors of a batch of nodes on the host's threads.  Long vectors (>= 2^12 superblocks) are
// decoded one after the other, each by several threads (decode_rrr63); the others side by side, one vector per thread —
// a tree of thousands of medium vectors loads at the host's aggregate decode rate, not at one core's.
inline void decode_nodes_parallel(const std::vector<FlatNode *> &batch) {
    if (batch.empty()) return;
    Binomial63::get();   // (built once, before any thread needs it)
    std::vector<FlatNode *> small;
    for (FlatNode *n : batch) {
        const uint64_t n_sb = ((n->raw.m_size + 63) / 63 + 31) / 32;
        if (n_sb >= (uint64_t(1) << 12) || batch.size() == 1) decode_rrr63(n->raw, &n->bits);
        else small.push_back(n);
    }
    if (small.empty()) return;
    unsigned n_thr = std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
    if (const char *e = std::getenv("GGQ_LOAD_THREADS")) n_thr = unsigned(std::max(1, std::atoi(e)));
    n_thr = unsigned(std::min<size_t>(n_thr, small.size()));
    if (n_thr <= 1) {
        for (FlatNode *n : small) decode_rrr63(n->raw, &n->bits, 1);
        return;
    }
    std::atomic<size_t> next{0};
    std::vector<std::string> errors(n_thr);
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < n_thr; ++t)
        pool.emplace_back([&, t] {
            try {
                for (size_t i = next.fetch_add(1); i < small.size(); i = next.fetch_add(1))
                    decode_rrr63(small[i]->raw, &small[i]->bits, 1);
            } catch (const std::exception &e) {
                errors[t] = e.what()[0] ? e.what() : "rrr: decode failed";
                next.store(small.size());   // the others stop at their next vector
            }
        });
    for (std::thread &th : pool) th.join();
    for (const std::string &e : errors)
        if (!e.empty()) throw ParseError(e);
}

// ---- step 1a: parsed pre-order tree -> BFS nodes ------------------------------------------
inline void flatten_parsed(ParsedNode &&root, FlatTree *t) {
    struct Pending { ParsedNode *p; int parent; int child_index; };
    t->nodes.clear();
    std::deque<Pending> q;
    q.push_back({&root, -1, 0});
    std::vector<ParsedNode *> src;
    while (!q.empty()) {
        Pending cur = q.front();
        q.pop_front();
        int id = int(t->nodes.size());
        t->nodes.emplace_back();
        src.push_back(cur.p);
        FlatNode &n = t->nodes.back();
        n.parent = cur.parent;
        n.child_index = cur.child_index;
        n.depth = cur.parent < 0 ? 0 : t->nodes[cur.parent].depth + 1;
        if (cur.parent >= 0) t->nodes[cur.parent].children.push_back(id);
        for (size_t i = 0; i < cur.p->children.size(); ++i)
            q.push_back({&cur.p->children[i], id, int(i)});
    }
    // columns: compose the partitions top-down
    for (size_t id = 0; id < t->nodes.size(); ++id) {
        FlatNode &n = t->nodes[id];
        ParsedNode &p = *src[id];
        if (n.parent < 0) {
            n.cols.resize(p.n_cols);
            for (uint32_t c = 0; c < p.n_cols; ++c) n.cols[c] = c;
        } else {
            const FlatNode &par = t->nodes[n.parent];
            const auto &map = src[n.parent]->partition[n.child_index];
            if (map.size() != p.n_cols) throw ParseError("BRWT: child column count does not match the parent's group");
            n.cols.resize(map.size());
            for (size_t c = 0; c < map.size(); ++c) n.cols[c] = par.cols[map[c]];
        }
        n.partition = p.partition;   // (kept for ggq_brwt_serialize; src[id]->partition is still read by the children)
        n.bits = std::move(p.bits);
        n.size = n.bits.size;
        n.ones = n.bits.ones;
        n.lazy = !p.raw.empty || (n.bits.w.empty() && n.bits.size != 0);
        n.raw = std::move(p.raw);
        if (n.lazy && !p.children.empty()) n.ones = p.children[0].bits.size;   // popcount = length of every child (checked at decode)
        if (n.parent >= 0 && n.size != t->nodes[n.parent].ones)
            throw ParseError("BRWT: child length differs from the parent's popcount");
    }
    t->n_rows = t->nodes[0].size;
    t->n_cols = t->nodes[0].cols.size();
}

// ---- step 1b: synthetic shape, basic partitioner with the given arity ----------------------
// Mirrors BRWTBottomUpBuilder::build with get_basic_partitioner(arity): consecutive runs of `arity`
// nodes are merged level by level; a run of one node is passed through unchanged
// (BRWT_builders.cpp:74-76), so leaves may sit at different depths.
inline void build_basic_shape(uint64_t n_rows, uint32_t n_cols, uint32_t arity, FlatTree *t) {
    struct Tmp { std::vector<int> children; std::vector<uint32_t> cols; };
    std::vector<Tmp> pool(n_cols);
    std::vector<int> level(n_cols);
    for (uint32_t j = 0; j < n_cols; ++j) { pool[j].cols = {j}; level[j] = int(j); }
    while (level.size() > 1) {
        std::vector<int> next;
        for (size_t i = 0; i < level.size(); i += arity) {
            size_t e = std::min(level.size(), i + arity);
            if (e - i == 1) { next.push_back(level[i]); continue; }
            Tmp p;
            for (size_t k = i; k < e; ++k) {
                p.children.push_back(level[k]);
                p.cols.insert(p.cols.end(), pool[level[k]].cols.begin(), pool[level[k]].cols.end());
            }
            pool.push_back(std::move(p));
            next.push_back(int(pool.size()) - 1);
        }
        level.swap(next);
    }
    t->nodes.clear();
    t->n_rows = n_rows;
    t->n_cols = n_cols;
    if (n_cols == 0) {  // BRWT(): a single empty node
        t->nodes.emplace_back();
        t->n_rows = 0;
        return;
    }
    struct Pending { int tmp; int parent; int child_index; };
    std::deque<Pending> q;
    q.push_back({level[0], -1, 0});
    while (!q.empty()) {
        Pending cur = q.front();
        q.pop_front();
        int id = int(t->nodes.size());
        t->nodes.emplace_back();
        FlatNode &n = t->nodes.back();
        n.parent = cur.parent;
        n.child_index = cur.child_index;
        n.depth = cur.parent < 0 ? 0 : t->nodes[cur.parent].depth + 1;
        n.cols = pool[cur.tmp].cols;
        if (cur.parent >= 0) t->nodes[cur.parent].children.push_back(id);
        for (size_t i = 0; i < pool[cur.tmp].children.size(); ++i)
            q.push_back({pool[cur.tmp].children[i], id, int(i)});
    }
    t->nodes[0].size = n_rows;
}

// ---- step 2: groups, lanes, DFS ranks, column paths (sizes may still be unknown) -----------
inline void assign_groups(FlatTree *t) {
    auto &N = t->nodes;
    t->group_hot.clear(); t->group_nblocks.clear(); t->group_parent.clear();
    t->lane_info.clear(); t->lane_aux.clear(); t->lane_node.clear();
    auto new_group = [&](uint32_t parent, uint32_t cnt, uint32_t more) {
        uint32_t sl = (cnt <= 1 ? 0 : (cnt == 2 ? 1 : 2)) + (t->slot_bits == kWideSlotBits ? 1 : 0);   // in 16-byte units
        t->group_hot.push_back(pack_group_hot(0, sl, cnt, more));
        t->group_nblocks.push_back(0);
        t->group_parent.push_back(parent);
        for (int k = 0; k < 4; ++k) { t->lane_info.push_back(0); t->lane_aux.push_back(0); t->lane_node.push_back(~0u); }
        return uint32_t(t->group_hot.size() - 1);
    };
    // group 0: the root alone
    uint32_t g0 = new_group(~0u, 1, 0);
    N[0].group = g0; N[0].lane = 0;
    t->lane_node[0] = 0;
    for (size_t id = 0; id < N.size(); ++id) {  // BFS order => a node's child groups are contiguous
        FlatNode &n = N[id];
        size_t a = n.children.size();
        if (!a) continue;
        n.first_child_group = uint32_t(t->group_hot.size());
        n.n_child_groups = uint32_t((a + kGroupLanes - 1) / kGroupLanes);
        for (size_t i = 0; i < a; i += kGroupLanes) {
            uint32_t cnt = uint32_t(std::min<size_t>(kGroupLanes, a - i));
            uint32_t g = new_group(uint32_t(id), cnt, i + kGroupLanes < a);
            for (uint32_t k = 0; k < cnt; ++k) {
                FlatNode &c = N[n.children[i + k]];
                c.group = g; c.lane = k;
                t->lane_node[g * 4 + k] = uint32_t(n.children[i + k]);
            }
        }
    }
    if (t->group_hot.size() > kMaxGroups) throw ParseError("BRWT has more sibling groups than the engine supports");
    // DFS leaf ranks
    t->leaf_col.clear();
    t->multi_column_leaves = false;
    t->max_depth = 0;
    std::vector<int> stack = {0};
    while (!stack.empty()) {
        int id = stack.back();
        stack.pop_back();
        FlatNode &n = N[id];
        t->max_depth = std::max<uint32_t>(t->max_depth, n.depth);
        if (n.children.empty()) {
            n.leaf_rank = uint32_t(t->leaf_col.size());
            for (uint32_t c : n.cols) t->leaf_col.push_back(c);
            if (n.cols.size() != 1) t->multi_column_leaves = true;
        } else {
            for (size_t i = n.children.size(); i-- > 0;) stack.push_back(n.children[i]);
        }
    }
    for (size_t id = 0; id < N.size(); ++id) {
        FlatNode &n = N[id];
        uint32_t li = n.group * 4 + n.lane;
        if (n.children.empty()) {
            t->lane_info[li] = kLeafFlag | n.leaf_rank;
            t->lane_aux[li] = uint32_t(n.cols.size());
        } else {
            t->lane_info[li] = n.first_child_group;
            t->lane_aux[li] = uint32_t(id);
        }
    }
    // column paths
    t->col_path_off.assign(t->n_cols + 1, 0);
    std::vector<int> leaf_of(t->n_cols, -1);
    for (size_t id = 0; id < N.size(); ++id)
        if (N[id].children.empty())
            for (uint32_t c : N[id].cols) leaf_of[c] = int(id);
    t->col_path.clear();
    for (uint64_t c = 0; c < t->n_cols; ++c) {
        std::vector<uint32_t> path;
        for (int id = leaf_of[c]; id >= 0; id = N[id].parent) path.push_back(N[id].group << 2 | N[id].lane);
        std::reverse(path.begin(), path.end());
        t->col_path_off[c] = uint32_t(t->col_path.size());
        t->col_path.insert(t->col_path.end(), path.begin(), path.end());
    }
    t->col_path_off[t->n_cols] = uint32_t(t->col_path.size());
}

// ---- step 2b: column sharding plan -----------------------------------------------------------
// Cut the tree at the smallest depth D whose frontier (nodes at depth D plus leaves above it, in DFS order) has
// at least 4 elements per shard; deal the frontier to shards in CONTIGUOUS DFS runs balanced by column count.
// Every rank derives the same plan from the tree alone.  Nodes at depth <= D stay on every shard; a deeper node
// is present only on the shard owning its depth-D ancestor.  Concatenating the shards' rows in shard order
// reproduces the reference's DFS output order (BRWT.cpp:45-50).
inline void plan_shards(FlatTree *t, uint32_t shard, uint32_t n_shards) {
    auto &N = t->nodes;
    t->shard = shard; t->n_shards = n_shards;
    t->col_owner.assign(t->n_cols, 0);
    if (n_shards <= 1) { t->node_present.clear(); t->col_owned.clear(); t->cut_depth = 0; return; }
    // DFS order of nodes
    std::vector<int> dfs;
    {
        std::vector<int> st = {0};
        while (!st.empty()) {
            int id = st.back(); st.pop_back();
            dfs.push_back(id);
            for (size_t i = N[id].children.size(); i-- > 0;) st.push_back(N[id].children[i]);
        }
    }
    uint32_t D = 1;
    std::vector<int> frontier;
    for (;; ++D) {
        frontier.clear();
        for (int id : dfs)
            if (uint32_t(N[id].depth) == D || (N[id].children.empty() && uint32_t(N[id].depth) < D)) frontier.push_back(id);
        if (frontier.size() >= size_t(4) * n_shards || D >= t->max_depth) break;
    }
    t->cut_depth = D;
    // owner of every frontier element: contiguous runs balanced by columns
    std::vector<int> owner_of_node(N.size(), -1);
    uint64_t before = 0, m = std::max<uint64_t>(t->n_cols, 1);
    for (int id : frontier) {
        uint64_t w = N[id].cols.size();
        uint64_t o = std::min<uint64_t>(n_shards - 1, (2 * before + w) * n_shards / (2 * m));
        owner_of_node[id] = int(o);
        before += w;
    }
    t->node_present.assign(N.size(), 0);
    t->col_owned.assign(t->n_cols, 0);
    for (size_t id = 0; id < N.size(); ++id) {   // BFS order: parents first
        FlatNode &n = N[id];
        if (owner_of_node[id] < 0 && uint32_t(n.depth) > D) owner_of_node[id] = owner_of_node[n.parent];
        bool above = uint32_t(n.depth) <= D;
        t->node_present[id] = above || owner_of_node[id] == int(shard);
        if (n.children.empty())
            for (uint32_t c : n.cols) { t->col_owner[c] = uint32_t(owner_of_node[id]); t->col_owned[c] = owner_of_node[id] == int(shard); }
        // frontier elements (and leaves above the cut) held by another shard: never descended / never reported
        if (owner_of_node[id] >= 0 && owner_of_node[id] != int(shard) && above)
            t->lane_info[n.group * 4 + n.lane] = kRemoteLane;
    }
}

// ---- step 3: arena geometry, once every node's size (= parent's ones) is known -------------
inline void assign_geometry(FlatTree *t) {
    auto &N = t->nodes;
    uint32_t G = t->n_groups();
    uint64_t slot = 0;
    t->lane_sb_off.assign(size_t(G) * 4, 0);
    t->lane_sel_off.assign(size_t(G) * 4, 0);
    uint64_t sb = 0, sel = 0;
    for (uint32_t g = 0; g < G; ++g) {
        uint64_t hot = t->group_hot[g];
        uint32_t sl = uint32_t(hot >> kGhStrideShift) & 3;
        uint32_t cnt = uint32_t(hot >> kGhCntShift) & 7;
        uint32_t more = uint32_t(hot >> kGhMoreShift) & 1;
        const bool g_present = t->present(t->lane_node[g * 4]);
        uint64_t len = g_present ? N[t->lane_node[g * 4]].size : 0;
        if (len >= kMaxRows) throw ParseError("vector longer than the engine supports");
        if (t->slot_bits == kWideSlotBits && len >= (uint64_t(1) << 32))
            throw ParseError("vector longer than the capacity layout supports (2^32 positions)");
        uint64_t nb = g_present ? len / t->slot_bits + 1 : 0;  // + sentinel block
        slot = (slot + 3) & ~uint64_t(3);   // 64-byte alignment of every group
        t->group_hot[g] = pack_group_hot(slot, sl, cnt, more) | (uint64_t(nb && ((nb - 1) >> kSuperShift) != 0) << kGhSuperShift);
        t->group_nblocks[g] = nb;
        slot += nb << sl;
        uint64_t n_sb = nb ? ((nb - 1) >> kSuperShift) + 1 : 1;
        for (uint32_t k = 0; k < cnt; ++k) {
            const FlatNode &n = N[t->lane_node[g * 4 + k]];
            t->lane_sb_off[g * 4 + k] = uint32_t(sb);
            sb += n_sb;
            t->lane_sel_off[g * 4 + k] = sel;
            if (g_present && (!n.children.empty() || n.parent < 0))  // select is only ever needed on internal nodes
                sel += ((n.ones + (uint64_t(1) << kSelectShift) - 1) >> kSelectShift) + 1;
        }
    }
    if (slot > kGhBaseMask) throw ParseError("arena larger than the engine supports");
    if (sb > 0xFFFFFFFFull) throw ParseError("too many superblocks");
    t->total_slots = (slot + 3) & ~uint64_t(3);
    t->total_sb = sb;
    t->total_sel = sel;
}

// ---- worst-case stack height of the descent kernel for this tree ---------------------------
// Replays the kernel's pop/push policy (ggq_engine.cu: brwt_rows_kernel) for `rows_per_warp` rows
// whose every bit is set — the row that pushes the most at every step.
inline uint32_t simulate_stack(const FlatTree &t, uint32_t rows_per_warp, uint32_t items_per_iter) {
    std::vector<uint32_t> st(rows_per_warp, 0);  // group ids; all rows start at the root group
    size_t max_top = st.size();
    std::vector<uint32_t> pushed, conts, kids0, kids1;
    while (!st.empty()) {
        size_t c = std::min<size_t>(st.size(), items_per_iter);
        pushed.clear();
        // the kernel handles the popped items in rounds of 16 (one item per lane pair and round) and pushes, per
        // round: continuation groups, then the first, then the second child of every lane
        for (size_t r0 = 0; r0 < c; r0 += 16) {
            conts.clear(); kids0.clear(); kids1.clear();
            for (size_t i = r0; i < std::min(c, r0 + 16); ++i) {
                uint32_t g = st[st.size() - 1 - i];
                uint64_t hot = t.group_hot[g];
                uint32_t cnt = uint32_t(hot >> kGhCntShift) & 7;
                if ((hot >> kGhMoreShift) & 1) conts.push_back(g + 1);
                for (uint32_t k = 0; k < cnt; ++k) {   // lane `half` pushes child 2*half first, then 2*half+1
                    uint32_t info = t.lane_info[g * 4 + k];
                    if (!(info & kLeafFlag)) ((k & 1) ? kids1 : kids0).push_back(info);
                }
            }
            pushed.insert(pushed.end(), conts.begin(), conts.end());
            pushed.insert(pushed.end(), kids0.begin(), kids0.end());
            pushed.insert(pushed.end(), kids1.begin(), kids1.end());
        }
        st.resize(st.size() - c);
        st.insert(st.end(), pushed.begin(), pushed.end());
        max_top = std::max(max_top, st.size());
    }
    return uint32_t(max_top);
}

}  // namespace ggq
