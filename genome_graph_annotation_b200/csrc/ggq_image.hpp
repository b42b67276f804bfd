// Flattened device-image file: the GPU-resident layout of a BRWT written next to the
// .brwt.annodbg it came from, so that a restart uploads the arrays as they are instead of re-parsing the
// reference stream (RRR decode, BRWT.cpp:87-111), re-flattening it and rebuilding rank headers / select samples.
//
//   header   magic "GGQIMG03", endianness tag, layout constants, matrix shape, the BRWT stats the reference API reports
//   tables   the host-side geometry (group records, lane tables, leaf order, column paths, shard plan) and the host tree
//            (every node's parent, children, column arrangement, vector length and popcount, RangePartition: what
//            ggq_brwt_serialize / ggq_tree_shape need) + FNV-1a
//   arrays   slots (16-byte slots with inline rank headers), select samples, superblock bases, lane totals, leaf codes
//   trailer  end tag (a truncated file is rejected)
//
// Host words are little-endian (sm_100a hosts are); the tag makes a foreign file fail loudly instead of misreading.
#pragma once
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace ggq {

static const char kImageMagic[8] = {'G', 'G', 'Q', 'I', 'M', 'G', '0', '3'};
static const u64 kImageEnd = 0x454e44474751494dull;
static const u32 kImageEndian = 0x01020304u;

struct ImageHeader {
    char magic[8];
    u32 endian, kind;
    u64 n_rows, n_cols, n_nodes, total_column_size, num_relations;
    u64 total_slots, total_sb, total_sel;
    u32 max_depth, multi_leaf, shard, n_shards, cut_depth, layout;   // layout: kImageLayout of the writing build
    u64 code_bytes, n_coded, code_escape_ppm, reserved2;
};
// the kernel constants an image depends on: a file written with other values must not be indexed with these
inline u32 image_layout(u32 slot_bits) { return slot_bits | (kSuperShift << 8) | (kSelectShift << 16) | (kGroupLanes << 24); }

struct Fnv {
    u64 h = 0xcbf29ce484222325ull;
    void add(const void *p, size_t n) {
        const unsigned char *b = static_cast<const unsigned char *>(p);
        for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 0x100000001b3ull; }
    }
};

struct ImageFile {
    FILE *f = nullptr;
    Fnv sum;
    explicit ImageFile(const char *path, const char *mode) : f(std::fopen(path, mode)) {
        if (!f) throw ParseError(std::string("cannot open image file ") + path);
    }
    ~ImageFile() { if (f) std::fclose(f); }
    void put(const void *p, size_t n) {
        if (n && std::fwrite(p, 1, n, f) != n) throw std::runtime_error("image write failed (disk full?)");
        sum.add(p, n);
    }
    void get(void *p, size_t n) {
        if (n && std::fread(p, 1, n, f) != n) throw ParseError("image file truncated");
        sum.add(p, n);
    }
    template <typename T> void put_vec(const std::vector<T> &v) {
        const u64 n = v.size();
        put(&n, 8);
        put(v.data(), size_t(n) * sizeof(T));
    }
    template <typename T> void get_vec(std::vector<T> *v, u64 limit) {
        u64 n = 0;
        get(&n, 8);
        if (n > limit) throw ParseError("image file: implausible table size");
        v->resize(size_t(n));
        get(v->data(), size_t(n) * sizeof(T));
    }
};

// device array <-> file through a pinned staging buffer (the arrays are tens of GB: no host copy of the whole)
template <typename T>
inline void image_put_device(ImageFile &io, const DeviceBuf<T> &d, unsigned char *stage, size_t stage_bytes) {
    const u64 bytes = d.n * sizeof(T);
    if (std::fwrite(&bytes, 1, 8, io.f) != 8) throw std::runtime_error("image write failed");
    const unsigned char *src = reinterpret_cast<const unsigned char *>(d.p);
    for (u64 o = 0; o < bytes; o += stage_bytes) {
        const size_t n = size_t(std::min<u64>(stage_bytes, bytes - o));
        CUDA_CHECK(cudaMemcpy(stage, src + o, n, cudaMemcpyDeviceToHost));
        if (std::fwrite(stage, 1, n, io.f) != n) throw std::runtime_error("image write failed (disk full?)");
    }
}

template <typename T>
inline void image_get_device(ImageFile &io, DeviceBuf<T> &d, u64 expect_count, unsigned char *stage, size_t stage_bytes) {
    u64 bytes = 0;
    if (std::fread(&bytes, 1, 8, io.f) != 8) throw ParseError("image file truncated");
    if (bytes != expect_count * sizeof(T)) throw ParseError("image file: array size does not match its header");
    d.alloc(size_t(expect_count));
    unsigned char *dst = reinterpret_cast<unsigned char *>(d.p);
    for (u64 o = 0; o < bytes; o += stage_bytes) {
        const size_t n = size_t(std::min<u64>(stage_bytes, bytes - o));
        if (std::fread(stage, 1, n, io.f) != n) throw ParseError("image file truncated");
        CUDA_CHECK(cudaMemcpy(dst + o, stage, n, cudaMemcpyHostToDevice));
    }
}

struct PinnedStage {
    unsigned char *p = nullptr;
    size_t bytes = size_t(64) << 20;
    PinnedStage() { CUDA_CHECK(cudaMallocHost(&p, bytes)); }
    ~PinnedStage() { if (p) cudaFreeHost(p); }
};

// The checksum proves the tables are the ones that were written, not that they are sane: a file from a damaged
// writer (or a forged one) must come back as GGQ_ERR_FORMAT, never as an out-of-bounds index on the host or device.
inline void image_check_tables(const FlatTree &t, const ImageHeader &h, const std::vector<u64> &lane_ones_h,
                               const std::vector<u64> &lane_size_h) {
    const u32 wide_shift = t.slot_bits == kWideSlotBits ? 1 : 0;   // stride_log2 counts 16-byte units
    const u64 G = t.group_hot.size();
    const u64 n_leaf_cols = t.leaf_col.size();
    auto bad = [](const char *what) { throw ParseError(std::string("image file: ") + what); };
    if (n_leaf_cols != h.n_cols && !(h.n_cols == 0 && n_leaf_cols == 0)) bad("leaf order does not cover the columns");
    for (u32 c : t.leaf_col) if (c >= h.n_cols) bad("leaf column out of range");
    for (u64 g = 0; g < G; ++g) {
        const u64 hot = t.group_hot[g];
        const u64 base = hot & kGhBaseMask, nb = t.group_nblocks[g];
        const u32 sl = u32(hot >> kGhStrideShift) & 3u, cnt = u32(hot >> kGhCntShift) & 7u;
        if (cnt > kGroupLanes || sl < wide_shift || (cnt > (1u << (sl - wide_shift)))) bad("group shape");
        if (wide_shift && nb > (u64(1) << 25)) bad("vector too long for the capacity layout");
        if (nb > (u64(1) << 40) || base + (nb << sl) > h.total_slots) bad("group outside the slot arena");
        if (((hot >> kGhMoreShift) & 1u) && g + 1 >= G) bad("continuation group missing");
        const u64 n_sb = nb ? ((nb - 1) >> kSuperShift) + 1 : 1;
        for (u32 k = 0; k < cnt; ++k) {
            const u64 li = g * 4 + k;
            const u32 info = t.lane_info[li];
            // every position of the vector lies inside the group's blocks
            if (lane_size_h[li] && (lane_ones_h[li] > lane_size_h[li] || lane_size_h[li] / t.slot_bits + 1 > nb)) bad("vector longer than its blocks");
            if (info == kRemoteLane) continue;
            if (info & kLeafFlag) {
                if (u64(info & ~kLeafFlag) + t.lane_aux[li] > n_leaf_cols) bad("leaf rank out of range");
            } else if (info >= G || info <= g) {
                bad("child group out of range");   // (level order: children come after their parent)
            }
            if (u64(t.lane_sb_off[li]) + n_sb > std::max<u64>(h.total_sb, 1)) bad("superblock bases out of range");
            if (!(info & kLeafFlag) || g == 0) {
                const u64 n_sel = ((lane_ones_h[li] + (u64(1) << kSelectShift) - 1) >> kSelectShift) + 1;
                if (nb && t.lane_sel_off[li] + n_sel > std::max<u64>(h.total_sel, 1)) bad("select samples out of range");
            }
        }
    }
    u32 prev = 0;
    for (u64 c = 0; c <= h.n_cols; ++c) {
        const u32 o = t.col_path_off[c];
        if (o < prev || o > t.col_path.size()) bad("column paths");
        prev = o;
    }
    for (u32 e : t.col_path) if ((e >> 2) >= G) bad("column path entry out of range");
    if (!t.lane_coded.empty() || !t.code_rec.empty() || h.code_bytes) {
        if (t.lane_coded.size() != 4 * G || t.code_rec.size() != 16 * G) bad("leaf-code tables");
        for (u64 li = 0; li < 4 * G; ++li) {
            if (!t.lane_coded[li]) continue;
            const u32 info = t.lane_info[li];
            if ((info & kLeafFlag) || info >= G) bad("coded lane is not internal");
            const u64 off = (u64(t.code_rec[4 * li + 1]) << 32) | t.code_rec[4 * li];
            if (off > h.code_bytes || lane_ones_h[li] > h.code_bytes - off) bad("leaf codes out of range");
            if (t.code_rec[4 * li + 3] > kCodeEscape || u64(t.code_rec[4 * li + 2]) + t.code_rec[4 * li + 3] > n_leaf_cols)
                bad("leaf-code ranks out of range");
        }
    }
}

// The host tree (FlatTree::nodes + lane_node), so that a matrix restored from an image is a full citizen: it can be
// serialized back into the reference's stream format and inspected node by node.
inline void image_put_nodes(ImageFile &io, const FlatTree &t) {
    const u64 n = t.nodes.size();
    io.put(&n, 8);
    for (const FlatNode &nd : t.nodes) {
        const int64_t fixed_i[1] = {nd.parent};
        const u32 fixed_u[7] = {u32(nd.child_index), u32(nd.depth), nd.group, nd.lane, nd.first_child_group, nd.n_child_groups, nd.leaf_rank};
        const u64 fixed_q[2] = {nd.size, nd.ones};
        io.put(fixed_i, sizeof(fixed_i)); io.put(fixed_u, sizeof(fixed_u)); io.put(fixed_q, sizeof(fixed_q));
        io.put_vec(nd.cols);
        io.put_vec(nd.children);
        const u64 ng = nd.partition.size();
        io.put(&ng, 8);
        for (const auto &g : nd.partition) io.put_vec(g);
    }
    io.put_vec(t.lane_node);
}

inline void image_get_nodes(ImageFile &io, FlatTree *t, u64 n_cols) {
    u64 n = 0;
    io.get(&n, 8);
    if (n > u64(kMaxGroups) * 4 + 1) throw ParseError("image file: implausible node count");
    t->nodes.assign(size_t(n), FlatNode());
    const u64 G = t->group_hot.size();
    for (u64 id = 0; id < n; ++id) {
        FlatNode &nd = t->nodes[id];
        int64_t fixed_i[1];
        u32 fixed_u[7];
        u64 fixed_q[2];
        io.get(fixed_i, sizeof(fixed_i)); io.get(fixed_u, sizeof(fixed_u)); io.get(fixed_q, sizeof(fixed_q));
        nd.parent = int(fixed_i[0]); nd.child_index = int(fixed_u[0]); nd.depth = int(fixed_u[1]);
        nd.group = fixed_u[2]; nd.lane = fixed_u[3]; nd.first_child_group = fixed_u[4]; nd.n_child_groups = fixed_u[5];
        nd.leaf_rank = fixed_u[6];
        nd.size = fixed_q[0]; nd.ones = fixed_q[1];
        io.get_vec(&nd.cols, n_cols);
        io.get_vec(&nd.children, n);
        u64 ng = 0;
        io.get(&ng, 8);
        if (ng > std::max<u64>(n_cols, 1)) throw ParseError("image file: implausible partition");
        nd.partition.assign(size_t(ng), {});
        for (auto &g : nd.partition) io.get_vec(&g, n_cols);
        // sanity: a node refers to earlier parents, later children, existing columns and an existing lane
        if ((id == 0) != (fixed_i[0] < 0) || fixed_i[0] >= int64_t(id) || nd.group >= G || nd.lane >= kGroupLanes ||
            nd.ones > nd.size || nd.size >= kMaxRows)
            throw ParseError("image file: host tree out of range");
        for (int c : nd.children)
            if (c <= int(id) || u64(c) >= n) throw ParseError("image file: host tree out of range");
        for (u32 c : nd.cols)
            if (c >= n_cols) throw ParseError("image file: host tree out of range");
        for (const auto &g : nd.partition)
            for (u32 c : g)
                if (c >= std::max<u64>(nd.cols.size(), 1)) throw ParseError("image file: host tree out of range");
    }
    io.get_vec(&t->lane_node, u64(kMaxGroups) * 4);
    if (n && t->lane_node.size() != 4 * G) throw ParseError("image file: inconsistent tables");
    for (u32 v : t->lane_node)
        if (v != ~0u && v >= n) throw ParseError("image file: host tree out of range");
}

inline void image_save(const Matrix *m, const char *path) {
    if (m->kind != Matrix::BRWT) throw std::runtime_error("only BRWT images can be saved");
    CUDA_CHECK(cudaSetDevice(m->device));
    const FlatTree &t = m->tree;
    ImageFile io(path, "wb");
    ImageHeader h;
    std::memset(&h, 0, sizeof(h));
    std::memcpy(h.magic, kImageMagic, 8);
    h.endian = kImageEndian; h.kind = 0;
    h.n_rows = t.n_rows; h.n_cols = t.n_cols;
    h.n_nodes = m->n_nodes; h.total_column_size = m->total_column_size; h.num_relations = m->num_relations;
    h.total_slots = m->slots.n; h.total_sb = m->sb_base.n; h.total_sel = m->sel.n;
    h.max_depth = t.max_depth; h.multi_leaf = t.multi_column_leaves; h.shard = t.shard; h.n_shards = t.n_shards;
    h.cut_depth = t.cut_depth; h.layout = image_layout(t.slot_bits);
    h.code_bytes = m->code_bytes; h.n_coded = m->n_coded; h.code_escape_ppm = u64(m->code_escape_rate * 1e6 + 0.5);
    io.put(&h, sizeof(h));
    io.put_vec(t.group_hot); io.put_vec(t.group_nblocks);
    io.put_vec(t.lane_info); io.put_vec(t.lane_aux); io.put_vec(t.lane_sb_off); io.put_vec(t.lane_sel_off);
    io.put_vec(t.leaf_col); io.put_vec(t.col_path_off); io.put_vec(t.col_path);
    io.put_vec(t.col_owned); io.put_vec(t.col_owner);
    io.put_vec(m->lane_size_h); io.put_vec(m->lane_ones_h);
    io.put_vec(t.lane_coded); io.put_vec(t.code_rec);
    image_put_nodes(io, t);
    const u64 check = io.sum.h;
    if (std::fwrite(&check, 1, 8, io.f) != 8) throw std::runtime_error("image write failed");
    PinnedStage st;
    image_put_device(io, m->slots, st.p, st.bytes);
    image_put_device(io, m->sel, st.p, st.bytes);
    image_put_device(io, m->sb_base, st.p, st.bytes);
    image_put_device(io, m->lane_ones, st.p, st.bytes);
    if (m->code_bytes) image_put_device(io, m->codes, st.p, st.bytes);
    if (std::fwrite(&kImageEnd, 1, 8, io.f) != 8 || std::fflush(io.f) != 0) throw std::runtime_error("image write failed");
}

inline Matrix *image_load(const char *path, int device) {
    ImageFile io(path, "rb");
    ImageHeader h;
    io.get(&h, sizeof(h));
    if (std::memcmp(h.magic, kImageMagic, 8) != 0) throw ParseError("not a ggq image file (or another version)");
    if (h.endian != kImageEndian || h.kind != 0) throw ParseError("image file of a foreign byte order / kind");
    if (h.layout != image_layout(kSlotBits) && h.layout != image_layout(kWideSlotBits))
        throw ParseError("image file written by a build with another slot layout");
    if (h.n_rows >= kMaxRows || h.n_cols >= (u64(1) << 31)) throw LimitError("image exceeds engine limits");
    std::unique_ptr<Matrix> m(new Matrix());
    m->kind = Matrix::BRWT;
    m->device = device;
    CUDA_CHECK(cudaSetDevice(device));
    FlatTree &t = m->tree;
    t.n_rows = h.n_rows; t.n_cols = h.n_cols;
    t.slot_bits = h.layout & 0xffu;
    const u64 lim = u64(1) << 33;
    io.get_vec(&t.group_hot, kMaxGroups); io.get_vec(&t.group_nblocks, kMaxGroups);
    io.get_vec(&t.lane_info, u64(kMaxGroups) * 4); io.get_vec(&t.lane_aux, u64(kMaxGroups) * 4);
    io.get_vec(&t.lane_sb_off, u64(kMaxGroups) * 4); io.get_vec(&t.lane_sel_off, u64(kMaxGroups) * 4);
    io.get_vec(&t.leaf_col, lim); io.get_vec(&t.col_path_off, lim); io.get_vec(&t.col_path, lim);
    io.get_vec(&t.col_owned, lim); io.get_vec(&t.col_owner, lim);
    io.get_vec(&m->lane_size_h, u64(kMaxGroups) * 4); io.get_vec(&m->lane_ones_h, u64(kMaxGroups) * 4);
    io.get_vec(&t.lane_coded, u64(kMaxGroups) * 4); io.get_vec(&t.code_rec, u64(kMaxGroups) * 16);
    image_get_nodes(io, &t, h.n_cols);
    const u64 expect = io.sum.h;
    u64 check = 0;
    if (std::fread(&check, 1, 8, io.f) != 8 || check != expect) throw ParseError("image file: table checksum mismatch");
    const size_t G = t.group_hot.size();
    if (G == 0 || t.group_nblocks.size() != G || t.lane_info.size() != 4 * G || t.lane_aux.size() != 4 * G ||
        t.lane_sb_off.size() != 4 * G || t.lane_sel_off.size() != 4 * G || m->lane_size_h.size() != 4 * G ||
        m->lane_ones_h.size() != 4 * G || t.col_path_off.size() != t.n_cols + 1 ||
        (!t.col_owned.empty() && t.col_owned.size() != t.n_cols))
        throw ParseError("image file: inconsistent tables");
    image_check_tables(t, h, m->lane_ones_h, m->lane_size_h);
    if (h.n_shards > 1) { t.nodes.clear(); t.lane_node.clear(); }   // (a shard's image does not carry which nodes it holds)
    t.total_slots = h.total_slots; t.total_sb = h.total_sb; t.total_sel = h.total_sel;
    t.max_depth = h.max_depth; t.multi_column_leaves = h.multi_leaf != 0;
    t.shard = h.shard; t.n_shards = h.n_shards; t.cut_depth = h.cut_depth;
    m->upload_tables(t);   // group records and lane tables; allocates lane_ones / sb_base
    PinnedStage st;
    image_get_device(io, m->slots, h.total_slots, st.p, st.bytes);
    image_get_device(io, m->sel, h.total_sel, st.p, st.bytes);
    image_get_device(io, m->sb_base, h.total_sb, st.p, st.bytes);
    image_get_device(io, m->lane_ones, u64(G) * 4, st.p, st.bytes);
    if (h.code_bytes) {
        image_get_device(io, m->codes, h.code_bytes, st.p, st.bytes);
        m->code_bytes = h.code_bytes; m->n_coded = u32(h.n_coded); m->code_escape_rate = double(h.code_escape_ppm) / 1e6;
        m->upload_codes_tables();
    }
    u64 end = 0;
    if (std::fread(&end, 1, 8, io.f) != 8 || end != kImageEnd) throw ParseError("image file truncated");
    m->refresh_img();
    m->n_nodes = h.n_nodes; m->total_column_size = h.total_column_size; m->num_relations = h.num_relations;
    m->derive_stats();   // no host nodes: keeps the stats above, configures the descent kernel
    return m.release();
}

}  // namespace ggq
