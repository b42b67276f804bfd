// K4: ColMajorCompressed (annotation/column_compressed/column_major.cpp:27-88) on the device: one
// Elias-Fano column (ggq_bv_sd.cuh) per matrix column, the SdImg descriptors in one device array.
//   get(row, col)   = bit_vector_sd::operator[]                      (column_major.cpp:27-32)
//   get_row(row)    = probe every column, ascending column ids        (column_major.cpp:34-47)
//   get_column(col) = bit_vector_sd::call_ones, ascending row ids     (column_major.cpp:49-56, bit_vector.cpp:767-790)
#pragma once
#include "ggq_bv.cuh"

namespace ggq {

struct ColMajorImage {
    std::vector<std::unique_ptr<SdBuf>> cols;
    std::vector<u64> col_ones;   // logical ones per column
    DeviceBuf<SdImg> d_cols;
    u64 n_rows = 0;
    size_t bytes() const {
        size_t b = d_cols.bytes();
        for (auto &c : cols) b += c->bytes();
        return b;
    }
};

__global__ void colmajor_get_kernel(const SdImg *cols, u32 n_cols, u64 n_rows, const u64 *rows, const u64 *qcols,
                                    u64 fixed_col, u64 n, unsigned char *out) {
    u64 t = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= n) return;
    u64 r = rows[t];
    u64 c = qcols ? qcols[t] : fixed_col;
    bool res = false;
    if (c < n_cols && r < n_rows) res = sd_get(cols[c], r);
    out[t] = res;
}

// one warp per (query row, 32 consecutive columns): lane j probes column 32*w + j
__global__ void colmajor_rows_probe_kernel(const SdImg *cols, u32 n_cols, u64 n_rows, const u64 *rows, u64 n,
                                           u32 words, u32 *bitmap, u32 *row_cnt) {
    u64 warp = (u64(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    u32 lane = threadIdx.x & 31;
    if (warp >= n * words) return;
    u64 t = warp / words;
    u32 w = u32(warp - t * words);
    u64 r = rows[t];
    u32 c = w * 32 + lane;
    bool bit = false;
    if (r < n_rows && c < n_cols) bit = sd_get(cols[c], r);
    u32 m = __ballot_sync(0xffffffffu, bit);
    if (lane == 0) {
        bitmap[warp] = m;
        if (m) atomicAdd(row_cnt + t, __popc(m));
    }
}

// one warp per query row: bitmap words -> ascending column ids at offsets[t]
template <typename OutT>
__global__ void colmajor_rows_emit_kernel(const u32 *bitmap, u32 words, const u64 *offsets, u64 n, OutT *out) {
    u64 t = (u64(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    u32 lane = threadIdx.x & 31;
    if (t >= n) return;
    u64 pos = offsets[t];
    for (u32 w0 = 0; w0 < words; w0 += 32) {
        u32 w = w0 + lane;
        u32 bits = w < words ? bitmap[t * words + w] : 0u;
        u32 c = __popc(bits), inc = c;
        for (int o = 1; o < 32; o <<= 1) {
            u32 v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        u64 p = pos + inc - c;
        while (bits) {
            u32 b = __ffs(bits) - 1;
            bits &= bits - 1;
            out[p++] = OutT(w * 32 + b);
        }
        pos += __shfl_sync(0xffffffffu, inc, 31);
    }
}

// non-inverted column: thread per block of the high part; out[k] = position of stored one #k
__global__ void sd_enumerate_kernel(SdImg v, u64 *out) {
    u64 b = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= v.high.n_blk) return;
    const u64 *blk = v.high.blk + b * kRbWords;
    u64 idx = blk[0];
    for (u32 j = 0; j < 7; ++j) {
        u64 x = blk[1 + j];
        while (x) {
            u32 t = __ffsll((long long)x) - 1;
            x &= x - 1;
            u64 hp = b * kRbBits + j * 64 + t;
            out[idx] = ((hp - idx) << v.wl) | sd_low(v, idx);
            ++idx;
        }
    }
}

// inverted column: thread per gap between consecutive stored ones (= zeros of the logical vector)
__global__ void sd_enumerate_inverted_kernel(SdImg v, u64 *out) {
    u64 k = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (k > v.m) return;
    u64 begin = k ? sd_stored_select(v, k - 1) + 1 : 0;
    u64 end = k < v.m ? sd_stored_select(v, k) : v.n;
    u64 o = begin - k;
    for (u64 p = begin; p < end; ++p) out[o++] = p;
}

inline Matrix *colmajor_load(Reader &in, int device) {
    CUDA_CHECK(cudaSetDevice(device));
    std::unique_ptr<Matrix> m(new Matrix());
    m->kind = Matrix::COLMAJOR;
    m->device = device;
    std::unique_ptr<ColMajorImage> cm(new ColMajorImage());
    u64 n_cols = in.u64be();
    if (n_cols > (1ull << 32)) throw ParseError("ColMajorCompressed: implausible column count");
    std::vector<SdImg> imgs;
    u64 relations = 0;
    for (u64 j = 0; j < n_cols; ++j) {
        SdColumn c;
        load_sd(in, &c);
        PlainBits bits;
        sd_to_plain(c, &bits);
        std::vector<u64> w = plain_to_words64(bits);
        DeviceBuf<u64> plain;
        plain.alloc(w.size());
        CUDA_CHECK(cudaMemcpy(plain.p, w.data(), w.size() * 8, cudaMemcpyHostToDevice));
        std::unique_ptr<SdBuf> sb(new SdBuf());
        sd_build(sb.get(), plain.p, bits.size, c.inverted ? 1 : 0, nullptr);
        if (j == 0) cm->n_rows = bits.size;
        imgs.push_back(sb->img);
        cm->col_ones.push_back(bits.ones);
        relations += bits.ones;
        cm->cols.push_back(std::move(sb));
    }
    cm->d_cols.alloc(std::max<size_t>(imgs.size(), 1));
    if (!imgs.empty()) CUDA_CHECK(cudaMemcpy(cm->d_cols.p, imgs.data(), imgs.size() * sizeof(SdImg), cudaMemcpyHostToDevice));
    m->tree.n_rows = cm->n_rows;
    m->tree.n_cols = n_cols;
    m->num_relations = relations;
    m->device_bytes = cm->bytes();
    CUDA_CHECK(cudaDeviceGetAttribute(&m->sm_count, cudaDevAttrMultiProcessorCount, device));
    m->cm = cm.release();
    return m.release();
}

inline void colmajor_rows_device(const Matrix *m, Workspace *ws, const u64 *d_rows, u64 n, int col_bytes, u64 *total_out) {
    const ColMajorImage *cm = m->cm;
    ws->reserve_rows(n);
    cudaStream_t st = ws->stream;
    u32 n_cols = u32(m->tree.n_cols);
    u32 words = std::max<u32>((n_cols + 31) / 32, 1);
    CUDA_CHECK(cudaMemsetAsync(ws->d_row_cnt.p, 0, (n + 1) * 4, st));
    if (ws->d_tmp.n < n * words) ws->d_tmp.alloc(n * words + 1024);
    ws->launches = 0;
    if (n) {
        u64 threads = n * words * 32;
        colmajor_rows_probe_kernel<<<u32((threads + 255) / 256), 256, 0, st>>>(cm->d_cols.p, n_cols, cm->n_rows, d_rows, n, words,
                                                                            ws->d_tmp.p, ws->d_row_cnt.p);
        CUDA_CHECK(cudaGetLastError());
        ws->launches += 1;
    }
    size_t cb = ws->cub_bytes;
    cub::DeviceScan::ExclusiveSum(ws->d_cub.p, cb, ws->d_row_cnt.p, ws->d_offsets.p, n + 1, st);
    ws->launches += 2;
    u64 total = 0;
    CUDA_CHECK(cudaMemcpyAsync(ws->h_ctrl, ws->d_offsets.p + n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    total = ws->h_ctrl[0];
    size_t out_bytes = size_t(total) * col_bytes;
    if (ws->d_out.n < out_bytes) ws->d_out.alloc(out_bytes + out_bytes / 8 + 1024);
    if (total) {
        u32 g = u32((n * 32 + 255) / 256);
        if (col_bytes == 8)
            colmajor_rows_emit_kernel<u64><<<g, 256, 0, st>>>(ws->d_tmp.p, words, ws->d_offsets.p, n, reinterpret_cast<u64 *>(ws->d_out.p));
        else
            colmajor_rows_emit_kernel<u32><<<g, 256, 0, st>>>(ws->d_tmp.p, words, ws->d_offsets.p, n, reinterpret_cast<u32 *>(ws->d_out.p));
        CUDA_CHECK(cudaGetLastError());
        ws->launches += 1;
    }
    ws->last_probes = 0;
    ws->last_touches = n * n_cols;
    ws->kernel_timed = false;
    *total_out = total;
}

inline void colmajor_get_launch(const Matrix *m, const u64 *d_rows, const u64 *d_cols, u64 fixed_col, u64 n,
                                unsigned char *d_out, cudaStream_t st) {
    const ColMajorImage *cm = m->cm;
    colmajor_get_kernel<<<u32((n + 255) / 256), 256, 0, st>>>(cm->d_cols.p, u32(m->tree.n_cols), cm->n_rows, d_rows, d_cols,
                                                              fixed_col, n, d_out);
}

inline void colmajor_get_column(const Matrix *m, Workspace *ws, u64 col, DeviceBuf<u64> *d_res, u64 *cnt) {
    const ColMajorImage *cm = m->cm;
    const SdImg &v = cm->cols[col]->img;
    *cnt = cm->col_ones[col];
    if (d_res->n < std::max<u64>(*cnt, 1)) d_res->alloc(std::max<u64>(*cnt, 1));
    if (!*cnt) return;
    if (!v.inverted)
        sd_enumerate_kernel<<<u32((v.high.n_blk + 127) / 128), 128, 0, ws->stream>>>(v, d_res->p);
    else
        sd_enumerate_inverted_kernel<<<u32((v.m + 1 + 127) / 128), 128, 0, ws->stream>>>(v, d_res->p);
    CUDA_CHECK(cudaGetLastError());
}

}  // namespace ggq
