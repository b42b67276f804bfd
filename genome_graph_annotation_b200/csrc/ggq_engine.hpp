// Host-side objects behind the opaque C handles.
#pragma once
#include <atomic>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "ggq_device.cuh"
#include "ggq_tree.hpp"

namespace ggq {

struct CudaError : std::runtime_error { using std::runtime_error::runtime_error; };
struct LimitError : std::runtime_error { using std::runtime_error::runtime_error; };
struct NcclError : std::runtime_error { using std::runtime_error::runtime_error; };

int fail(int code, const std::string &msg);
u32 default_slot_bits();   // slot layout of the matrices created from now on (ggq_set_default_option)

#define CUDA_CHECK(expr)                                                                          \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess) {                                                                  \
            cudaGetLastError(); /* clear the thread's last-error slot: a later launch check must not see this one */ \
            if (_e == cudaErrorMemoryAllocation) throw std::bad_alloc();                          \
            throw ggq::CudaError(std::string(#expr) + ": " + cudaGetErrorString(_e));             \
        }                                                                                         \
    } while (0)

template <typename T>
struct DeviceBuf {
    T *p = nullptr;
    size_t n = 0;
    DeviceBuf() = default;
    DeviceBuf(const DeviceBuf &) = delete;
    DeviceBuf &operator=(const DeviceBuf &) = delete;
    DeviceBuf(DeviceBuf &&o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    DeviceBuf &operator=(DeviceBuf &&o) noexcept {
        if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; }
        return *this;
    }
    ~DeviceBuf() { release(); }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    void alloc(size_t count) {
        release();
        if (!count) count = 1;
        CUDA_CHECK(cudaMalloc(reinterpret_cast<void **>(&p), count * sizeof(T)));
        n = count;
    }
    size_t bytes() const { return n * sizeof(T); }
};

struct Workspace;
struct ColMajorImage;

struct Matrix {
    enum Kind { BRWT = 0, COLMAJOR = 1 } kind = BRWT;
    int device = 0;
    int sm_count = 148;
    FlatTree tree;   // host copy of the structure (bits dropped after upload)

    // device image
    DeviceBuf<uint4> slots;
    DeviceBuf<u64> group_hot, group_nblocks, lane_sel_off, lane_ones, sb_base;
    DeviceBuf<uint4> group_rec;   // per (group, half): hot word + the two children's lane_info (descent kernel)
    DeviceBuf<uint4> group_rec16; // per group: compact record (trees with < 2^20 - 1 columns), else empty
    bool compact_rec = false;
    DeviceBuf<u32> lane_info, lane_aux, lane_sb_off, sel, leaf_col, col_path_off, col_path;
    // leaf codes (ggq_rows.cuh): one byte per position of the coded nodes' child vectors
    DeviceBuf<unsigned char> codes;
    DeviceBuf<uint4> code_rec;
    bool use_codes = true;        // ggq_set_option("leaf_codes", 0/1)
    bool wide() const { return tree.slot_bits == kWideSlotBits; }   // capacity layout (ggq_layout.hpp)
    u64 code_bytes = 0;
    u32 n_coded = 0;
    double code_escape_rate = 0;  // visit-weighted escape fraction of the coded nodes (sampled at build time)
    Img img{};

    // stats
    u64 n_nodes = 0, total_column_size = 0, num_relations = 0, device_bytes = 0;
    mutable std::atomic<u64> last_probes{0};   // const calls from many threads store it (relaxed)
    std::vector<u64> lane_size_h, lane_ones_h;   // per lane: bits / set bits of its vector (host copy)

    // descent kernel configuration
    u32 bm_words = 1, rb = 8, cap_s = 64, cap_spill = 0, rows_grid = 148;
    // cap_s / cap_spill come from a replay of the kernel's pop/push policy (simulate_stack): an estimate.  The kernel
    // never writes past it; when a batch would, it raises a flag and the engine switches the matrix to the spill
    // instantiation with cap_safe = rows per tile x leaves of the group tree (pending items of one row are disjoint
    // subtrees of the group tree, so this many can never be exceeded) and reruns the batch.
    u32 cap_safe = 0;
    mutable std::atomic<bool> stack_safe{false};
    u32 root_child = 0xFFFFFFFFu, root_sl = 0, root_multi_sb = 0, any_more = 0, root_all_ones = 0;   // K1 root fast path (RowsArgs)
    u64 root_base = 0;
    size_t rows_smem = 0;
    // pipelined host calls: one CTA per SM fewer (shared-memory request padded so that no more fit), which leaves
    // room for the small kernels of the neighbouring chunks (ordering, scan, CSR gather) beside a running descent
    u32 rows_grid_p = 148;
    size_t rows_smem_p = 0;
    u32 sort_bits = 0, row_bits = 0;   // query ordering: sort on the top sort_bits of the row_bits-bit row id
    u64 sort_min_n = 4096;             // batches smaller than this are processed in the caller's order

    ColMajorImage *cm = nullptr;  // column-major matrices

    mutable std::mutex pool_mu;
    mutable std::vector<Workspace *> pool;

    Matrix() = default;
    ~Matrix();
    void upload_tables(const FlatTree &t);
    void upload_records(const FlatTree &t);
    void build_leaf_codes();
    void upload_codes_tables();
    void finalize_arena(const FlatTree &t, bool ones_known);
    void refresh_img();
    void derive_stats();
    void configure_rows(u32 rb_request, u32 sort_request);
    Workspace *acquire() const;
    void release(Workspace *w) const;
};

// Per-caller scratch: one stream, device buffers that grow to the high-water mark, pinned staging.
struct Workspace {
    const Matrix *m;
    cudaStream_t stream = nullptr;
    DeviceBuf<u64> d_rows, d_row_tmp_off, d_offsets, d_bytes8, d_ctrl, d_spill;
    DeviceBuf<u32> d_row_cnt, d_tmp, d_iota, d_perm, d_rows32;   // d_rows32: u32 row ids of the compact host call
    DeviceBuf<unsigned char> d_cnt_out;                          // narrowed per-row counts of the compact host call
    DeviceBuf<u64> d_rows_sorted;
    DeviceBuf<unsigned char> d_sort_tmp;
    size_t sort_tmp_bytes = 0;
    DeviceBuf<int> d_err;
    DeviceBuf<unsigned char> d_out, d_cub;
    size_t cub_bytes = 0;
    u64 cap_rows = 0;
    bool ctrl_clean = false;       // d_ctrl[0..8] are zero: the last direct label call cleared them on the device (no memset needed)
    u64 *h_ctrl = nullptr;
    unsigned char *h_stage = nullptr;
    size_t h_stage_bytes = 0;
    u64 last_probes = 0, last_touches = 0, last_code_probes = 0, last_code_escapes = 0;
    int launches = 0;
    int pending_sort_launches = 0;   // between brwt_rows_begin and brwt_rows_finish
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;   // around the dominant kernel of the last query
    bool kernel_timed = false;
    bool ran_safe = false;           // last descent ran with the provable stack capacity (Matrix::stack_safe)
    bool pipelined = false;          // chunk of a pipelined host call: counting-sort ordering, descent leaves a CTA slot free
    DeviceBuf<u32> d_order;          // bucket counters of the counting sort
    DeviceBuf<u64> d_blk;            // block sums of the chunk's offset scan
    bool hist_mode = false;          // count_labels: skip CSR assembly, the ids stay unordered in d_tmp
    DeviceBuf<u64> d_counts;
    DeviceBuf<u64> d_col;            // get_column result
    DeviceBuf<u64> d_lab_cols, d_lab_cnt, d_lab_cols2, d_lab_cnt2, d_lab_n;   // get_labels / get_top_labels
    DeviceBuf<unsigned char> d_lab_sort;
    DeviceBuf<u64> d_qoff, d_qlab_off, d_qout_cols, d_qout_cnt;   // ggq_get_labels_batch
    DeviceBuf<u32> d_qstage_cols, d_qstage_cnt, d_qkept;
    // column-sharded queries: gathered partial CSRs of all shards, merged result, exchange timing
    DeviceBuf<u64> g_off, g_tot, m_off;
    DeviceBuf<u32> g_cols;
    DeviceBuf<unsigned char> m_out;
    cudaEvent_t ev2 = nullptr, ev3 = nullptr;
    cudaEvent_t ev_ctrl = nullptr;   // pipelined chunks: control words published (before the offset scan)
    bool exchange_timed = false;
    u64 exchange_bytes = 0;

    explicit Workspace(const Matrix *m_);
    ~Workspace();
    void reserve_rows(u64 n);
    void stage_reserve(size_t bytes);
};

void build_chunk_table(const FlatTree &t, std::vector<u32> *chunk_lane, std::vector<u32> *chunk_first,
                       std::vector<u32> *lane_chunk_off);

}  // namespace ggq
