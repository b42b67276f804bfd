// Column-sharded row queries (SURVEY.md 8e "root-subtree column sharding", config C5).
//
// Every rank holds the replicated top of the tree and the subtrees of its own run of the cut frontier
// (ggq_tree.hpp: plan_shards).  A query batch is answered in three steps:
//   1. every rank runs K1 on the FULL batch over its own subtrees  -> partial CSR (u64 offsets, u32 column ids);
//   2. exchange: one ncclAllGather of the partial offsets and one grouped ncclBroadcast per rank of the partial
//      column ids (an all-gather-v: no padding crosses NVLink);
//   3. merge_count_kernel + exclusive scan + merge_copy_kernel concatenate, per query, the shards' segments in
//      SHARD ORDER — shards own contiguous DFS runs of the frontier, so this is the reference's output order
//      (BRWT.cpp:45-50) with no sort.
// ggq_merge_shards_device is step 3 alone on caller-provided pointers: it is what a single-GPU test (all shards
// on one device), a host that runs its own collective, or a peer-memory (cudaIpc / NVLink P2P) variant calls.
#pragma once
#include <dlfcn.h>
#include <nccl.h>   // types only: the library is bound at first use (see NcclApi)

#include "ggq_engine.hpp"

namespace ggq {

constexpr int kMaxShards = 16;

struct ShardPtrs {
    const u64 *off[kMaxShards];
    const u32 *cols[kMaxShards];
    int n;
};

__global__ void merge_count_kernel(ShardPtrs p, u64 n, u32 *cnt) {
    u64 q = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (q >= n) return;
    u64 c = 0;
    for (int r = 0; r < p.n; ++r) c += p.off[r][q + 1] - p.off[r][q];
    cnt[q] = u32(c);
}

// 8 lanes per query
template <typename OutT>
__global__ void merge_copy_kernel(ShardPtrs p, const u64 *offsets, u64 n, OutT *out) {
    const u64 q = (u64(blockIdx.x) * blockDim.x + threadIdx.x) >> 3;
    const u32 sub = threadIdx.x & 7;
    if (q >= n) return;
    u64 dst = offsets[q];
    for (int r = 0; r < p.n; ++r) {
        const u64 b = p.off[r][q], len = p.off[r][q + 1] - b;
        const u32 *src = p.cols[r] + b;
        for (u64 i = sub; i < len; i += 8) out[dst + i] = OutT(src[i]);
        dst += len;
    }
}

struct Comm {
    ncclComm_t comm = nullptr;
    int rank = 0, n_ranks = 1, device = 0;
    // Peer-memory exchange (ggq_comm_enable_peer): the partial-CSR buffers of one workspace per rank are exported
    // with cudaIpc and mapped by every other rank, so the merge kernels read the shards' segments straight over
    // NVLink/NVSwitch (P2P loads) — no all-gather, no staging copy.  Valid while batches stay within the capacities
    // the buffers were pinned at; a larger batch falls back to the NCCL exchange for good (all ranks decide alike).
    bool peer = false;
    Workspace *peer_ws = nullptr;        // the local workspace whose buffers were exported
    u64 peer_max_rows = 0, peer_max_ids = 0;
    const u64 *peer_off[16] = {};        // [r]: rank r's partial offsets (null for the own rank)
    const u32 *peer_cols[16] = {};       // [r]: rank r's partial column ids
    void close_peer() {
        for (int r = 0; r < 16; ++r) {
            if (peer_off[r]) cudaIpcCloseMemHandle(const_cast<u64 *>(peer_off[r]));
            if (peer_cols[r]) cudaIpcCloseMemHandle(const_cast<u32 *>(peer_cols[r]));
            peer_off[r] = nullptr;
            peer_cols[r] = nullptr;
        }
        peer = false;
        peer_ws = nullptr;
    }
    ~Comm() { close_peer(); }
};

// NCCL is resolved with dlopen at the first collective call instead of being a link-time dependency: a host
// process that already carries an NCCL (e.g. the one bundled with PyTorch) keeps using that one, and hosts that
// never shard need no NCCL at all.
struct NcclApi {
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclBroadcast) Broadcast = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    static NcclApi &get() {
        static NcclApi api = load();
        return api;
    }
    static NcclApi load() {
        NcclApi a;
        void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) throw NcclError(std::string("cannot load libnccl.so.2: ") + dlerror());
        auto sym = [&](const char *name) {
            void *p = dlsym(h, name);
            if (!p) throw NcclError(std::string("libnccl lacks ") + name);
            return p;
        };
        a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
        a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
        a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
        a.AllGather = reinterpret_cast<decltype(a.AllGather)>(sym("ncclAllGather"));
        a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(sym("ncclAllReduce"));
        a.Broadcast = reinterpret_cast<decltype(a.Broadcast)>(sym("ncclBroadcast"));
        a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
        a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
        a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
        return a;
    }
};

#define NCCL_CHECK(expr)                                                                         \
    do {                                                                                         \
        ncclResult_t _r = (expr);                                                                \
        if (_r != ncclSuccess) throw ggq::NcclError(std::string(#expr) + ": " + ggq::NcclApi::get().GetErrorString(_r)); \
    } while (0)

// step 3 on the workspace stream; result in ws->m_off / ws->m_out
inline void merge_shards(Workspace *ws, const ShardPtrs &p, u64 n, int col_bytes, u64 *total_out) {
    cudaStream_t st = ws->stream;
    ws->reserve_rows(n);
    if (ws->m_off.n < n + 1) ws->m_off.alloc(std::max<u64>(n + 1, ws->cap_rows));
    if (n == 0) {
        CUDA_CHECK(cudaMemsetAsync(ws->m_off.p, 0, 8, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        *total_out = 0;
        return;
    }
    merge_count_kernel<<<u32((n + 255) / 256), 256, 0, st>>>(p, n, ws->d_row_cnt.p);
    CUDA_CHECK(cudaMemsetAsync(ws->d_row_cnt.p + n, 0, 4, st));
    size_t cb = ws->cub_bytes;
    cub::DeviceScan::ExclusiveSum(ws->d_cub.p, cb, ws->d_row_cnt.p, ws->m_off.p, n + 1, st);
    CUDA_CHECK(cudaMemcpyAsync(ws->h_ctrl, ws->m_off.p + n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    const u64 total = ws->h_ctrl[0];
    const size_t out_bytes = size_t(total) * col_bytes;
    if (ws->m_out.n < out_bytes) ws->m_out.alloc(out_bytes + out_bytes / 8 + 1024);
    if (total) {
        const u32 g = u32((n * 8 + 255) / 256);
        if (col_bytes == 8) merge_copy_kernel<u64><<<g, 256, 0, st>>>(p, ws->m_off.p, n, reinterpret_cast<u64 *>(ws->m_out.p));
        else merge_copy_kernel<u32><<<g, 256, 0, st>>>(p, ws->m_off.p, n, reinterpret_cast<u32 *>(ws->m_out.p));
        CUDA_CHECK(cudaGetLastError());
    }
    *total_out = total;
}

void rows_device(const Matrix *m, Workspace *ws, const u64 *d_rows, u64 n, int col_bytes, u64 *total_out);

// steps 1-3.  slice == false: every rank ends with the complete CSR of the batch.  slice == true: rank k merges only
// ITS slice of the batch, queries [n*k/R, n*(k+1)/R) — the consumer of a sharded query is itself spread over the
// ranks, and merging everything everywhere makes the exchange (R-1)/R of the whole result per rank, which does not
// shrink with R.  With the peer-memory exchange a slice reads 1/R of that over NVLink; the NCCL exchange still
// gathers everything (correct, not faster).
inline void rows_sharded_device(const Matrix *m, Workspace *ws, Comm *c, const u64 *d_rows, u64 n, int col_bytes,
                                u64 *total_out, bool slice = false, u64 *slice_first = nullptr, u64 *slice_n = nullptr) {
    const u64 q0 = slice ? n * u64(c->rank) / u64(c->n_ranks) : 0;
    const u64 q1 = slice ? n * u64(c->rank + 1) / u64(c->n_ranks) : n;
    if (slice_first) *slice_first = q0;
    if (slice_n) *slice_n = q1 - q0;
    if (c->n_ranks > kMaxShards) throw LimitError("too many shards");
    if (u32(c->n_ranks) != m->tree.n_shards || u32(c->rank) != m->tree.shard)
        throw std::runtime_error("communicator rank/size does not match the matrix shard");
    cudaStream_t st = ws->stream;
    const int R = c->n_ranks;
    u64 total = 0;
    rows_device(m, ws, d_rows, n, 4, &total);                      // 1. partial CSR in ws->d_offsets / ws->d_out
    cudaEvent_t &e0 = ws->ev2, &e1 = ws->ev3;
    CUDA_CHECK(cudaEventRecord(e0, st));
    // 2a. totals (8 bytes per rank) -> sizes of the column-id segments
    if (ws->g_tot.n < size_t(kMaxShards) + 1) ws->g_tot.alloc(kMaxShards + 1);
    CUDA_CHECK(cudaMemcpyAsync(ws->g_tot.p + kMaxShards, ws->d_offsets.p + n, 8, cudaMemcpyDeviceToDevice, st));
    NCCL_CHECK(NcclApi::get().AllGather(ws->g_tot.p + kMaxShards, ws->g_tot.p, 1, ncclUint64, c->comm, st));
    CUDA_CHECK(cudaMemcpyAsync(ws->h_ctrl, ws->g_tot.p, size_t(R) * 8, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    u64 tot[kMaxShards], col_off[kMaxShards + 1];
    col_off[0] = 0;
    for (int r = 0; r < R; ++r) { tot[r] = ws->h_ctrl[r]; col_off[r + 1] = col_off[r] + tot[r]; }
    if (c->peer && ws == c->peer_ws) {
        bool within = n <= c->peer_max_rows;
        for (int r = 0; r < R; ++r) within = within && tot[r] <= c->peer_max_ids;
        if (within) {
            // 2'. no transfer step: the all-gather of the totals above is the barrier ("every shard's partial CSR is
            // complete"); the merge kernels gather the segments from peer memory, tile by tile
            ShardPtrs p;
            p.n = R;
            for (int r = 0; r < R; ++r) {
                p.off[r] = (r == c->rank ? ws->d_offsets.p : c->peer_off[r]) + q0;   // (offsets are absolute: cols stay)
                p.cols[r] = r == c->rank ? reinterpret_cast<const u32 *>(ws->d_out.p) : c->peer_cols[r];
            }
            merge_shards(ws, p, q1 - q0, col_bytes, total_out);
            // nobody may overwrite its partial CSR (next query) before every peer has read it
            NCCL_CHECK(NcclApi::get().AllGather(ws->g_tot.p + kMaxShards, ws->g_tot.p, 1, ncclUint64, c->comm, st));
            CUDA_CHECK(cudaEventRecord(e1, st));
            ws->exchange_bytes = 0;
            for (int r = 0; r < R; ++r)   // offsets are read by both merge kernels; ids of the merged queries once
                if (r != c->rank) ws->exchange_bytes += (q1 - q0 + 1) * 8 * 2 + (slice ? tot[r] / u64(R) : tot[r]) * 4;
            ws->exchange_timed = true;
            ws->launches += 5;
            return;
        }
        c->peer = false;   // the exported buffers were (or will be) reallocated: the mappings are stale
    }
    if (ws->g_off.n < size_t(R) * (n + 1)) ws->g_off.alloc(size_t(R) * (n + 1));
    if (ws->g_cols.n < std::max<u64>(col_off[R], 1)) ws->g_cols.alloc(col_off[R] + col_off[R] / 8 + 1024);
    // 2b. offsets: all-gather; column ids: all-gather-v as grouped broadcasts
    NCCL_CHECK(NcclApi::get().AllGather(ws->d_offsets.p, ws->g_off.p, n + 1, ncclUint64, c->comm, st));
    NCCL_CHECK(NcclApi::get().GroupStart());
    for (int r = 0; r < R; ++r)
        if (tot[r])
            NCCL_CHECK(NcclApi::get().Broadcast(ws->d_out.p, ws->g_cols.p + col_off[r], tot[r], ncclUint32, r, c->comm, st));
    NCCL_CHECK(NcclApi::get().GroupEnd());
    CUDA_CHECK(cudaEventRecord(e1, st));
    ws->exchange_bytes = u64(R) * (n + 1) * 8 + col_off[R] * 4;
    // 3. merge in shard order
    ShardPtrs p;
    p.n = R;
    for (int r = 0; r < R; ++r) { p.off[r] = ws->g_off.p + size_t(r) * (n + 1) + q0; p.cols[r] = ws->g_cols.p + col_off[r]; }
    merge_shards(ws, p, q1 - q0, col_bytes, total_out);
    ws->exchange_timed = true;
    ws->launches += 3 + R;
}

}  // namespace ggq
