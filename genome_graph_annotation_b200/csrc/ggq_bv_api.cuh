// Host object + query kernels + C ABI of the bit vectors (include/ggq.h "bit vectors").
#pragma once

namespace ggq {

enum BvOp { OP_RANK1 = 0, OP_SELECT1 = 1, OP_GET = 2, OP_RANK0 = 3, OP_SELECT0 = 4, OP_GETINT = 5, OP_WORDS = 6, OP_MAPRANK = 7 };

struct BvImg {
    int kind;
    u64 n_bits, n_ones;
    RbImg stat;
    SdImg sd;
    RrrImg rrr;
};

// One thread per query, grid-stride.  OutT = u64 (rank/select) or unsigned char (get).
template <int KIND, int OP, typename OutT>
__global__ void __launch_bounds__(256) bv_query_kernel(BvImg v, const u64 *q, u64 n, OutT *out) {
    __shared__ RrrTables T;
    if (KIND == GGQ_BV_RRR) {
        const u64 *src = reinterpret_cast<const u64 *>(&g_rrr_tables);
        u64 *dst = reinterpret_cast<u64 *>(&T);
        for (u32 i = threadIdx.x; i < sizeof(RrrTables) / 8; i += blockDim.x) dst[i] = src[i];
        __syncthreads();
    }
    for (u64 t = u64(blockIdx.x) * blockDim.x + threadIdx.x; t < n; t += u64(gridDim.x) * blockDim.x) {
        u64 x = q[t];
        u64 res = 0;
        if (OP == OP_SELECT1) {
            // k in [1, n_ones]; anything else is undefined in the reference (assert only): return n_bits
            if (x < 1 || x > v.n_ones) res = v.n_bits;
            else if (KIND == GGQ_BV_STAT) res = rb_select<false>(v.stat, x - 1);
            else if (KIND == GGQ_BV_SD) res = sd_select1(v.sd, x);
            else res = rrr_select(v.rrr, T, x - 1);
        } else if (OP == OP_RANK1) {
            if (v.n_bits == 0) res = 0;
            else {
                u64 i = x < v.n_bits ? x : v.n_bits - 1;   // clamp (bit_vector.cpp:285,692,860)
                bool bit;
                if (KIND == GGQ_BV_STAT) { res = rb_rank_excl_bit(v.stat, i, &bit); res += bit; }
                else if (KIND == GGQ_BV_SD) res = sd_rank1(v.sd, i);
                else res = rrr_rank_incl_bit(v.rrr, T, i, &bit);
            }
        } else {
            bool bit = false;
            if (x < v.n_bits) {
                if (KIND == GGQ_BV_STAT) rb_rank_excl_bit(v.stat, x, &bit);
                else if (KIND == GGQ_BV_SD) bit = sd_get(v.sd, x);
                else rrr_rank_incl_bit(v.rrr, T, x, &bit);
            }
            res = bit;
        }
        out[t] = OutT(res);
    }
}

// The rest of the reference's bit_vector interface, by composition of rank1 / operator[] (none of it is on the row
// query path; the reference composes them the same way):
//   rank0(id)  = min(id + 1, size) - rank1(id)                                  (common/bit_vector.cpp:32-34)
//   select0(k) = smallest i with rank0(i) = k, k in [1, size - num_set_bits]     (bit_vector_rrr::select0, bit_vector.hpp:199;
//                by bisection over rank1 — every kind answers it, the reference offers it on rrr only)
//   get_int(id, width) = bits [id, id + width) as an integer, width <= 64        (bit_vector.hpp:25)
//   words      = the plain bits, 64 per thread (to_vector / the sdsl::bit_vector copy, bit_vector.cpp:36-46)
template <int KIND>
__device__ __forceinline__ u64 bv_rank1_incl(const BvImg &v, const RrrTables &T, u64 i, bool *bit) {   // i < n_bits
    if (KIND == GGQ_BV_STAT) { u64 r = rb_rank_excl_bit(v.stat, i, bit); return r + *bit; }
    if (KIND == GGQ_BV_SD) { *bit = sd_get(v.sd, i); return sd_rank1(v.sd, i); }
    return rrr_rank_incl_bit(v.rrr, T, i, bit);
}

template <int KIND, int OP>
__global__ void __launch_bounds__(256) bv_aux_kernel(BvImg v, const u64 *q, const u32 *widths, u32 width, u64 n, u64 *out) {
    __shared__ RrrTables T;
    if (KIND == GGQ_BV_RRR) {
        const u64 *src = reinterpret_cast<const u64 *>(&g_rrr_tables);
        u64 *dst = reinterpret_cast<u64 *>(&T);
        for (u32 i = threadIdx.x; i < sizeof(RrrTables) / 8; i += blockDim.x) dst[i] = src[i];
        __syncthreads();
    }
    for (u64 t = u64(blockIdx.x) * blockDim.x + threadIdx.x; t < n; t += u64(gridDim.x) * blockDim.x) {
        bool bit;
        u64 res = 0;
        if (OP == OP_RANK0) {
            const u64 x = q[t];
            if (v.n_bits) {
                const u64 i = x < v.n_bits ? x : v.n_bits - 1;
                res = i + 1 - bv_rank1_incl<KIND>(v, T, i, &bit);
            }
        } else if (OP == OP_SELECT0) {
            const u64 k = q[t];
            res = v.n_bits;   // outside [1, zeros]: undefined in the reference (assert only)
            if (k >= 1 && k <= v.n_bits - v.n_ones) {
                u64 lo = 0, hi = v.n_bits - 1;   // smallest i with rank0(i) >= k
                while (lo < hi) {
                    const u64 mid = lo + ((hi - lo) >> 1);
                    if (mid + 1 - bv_rank1_incl<KIND>(v, T, mid, &bit) >= k) hi = mid; else lo = mid + 1;
                }
                res = lo;
            }
        } else if (OP == OP_GETINT) {
            const u64 x = q[t];
            const u32 w = widths ? widths[t] : width;
            for (u32 j = 0; j < w && j < 64 && x + j < v.n_bits; ++j) {
                bv_rank1_incl<KIND>(v, T, x + j, &bit);
                res |= u64(bit) << j;
            }
        } else if (OP == OP_MAPRANK) {
            // AnnotatedDBG::graph_to_anno_index (common/annotated_dbg.cpp:142-156): rank1(i) - 1 where the mask is set;
            // the reference throws where it is not — here such an index maps to UINT64_MAX, which every row query
            // treats as an empty row
            const u64 x = q[t];
            res = ~0ull;
            if (x < v.n_bits) {
                const u64 r = bv_rank1_incl<KIND>(v, T, x, &bit);
                if (bit) res = r - 1;
            }
        } else {   // OP_WORDS: word t of the plain bits
            for (u32 j = 0; j < 64 && t * 64 + j < v.n_bits; ++j) {
                bv_rank1_incl<KIND>(v, T, t * 64 + j, &bit);
                res |= u64(bit) << j;
            }
        }
        out[t] = res;
    }
}

__global__ void iota1_kernel(u64 *p, u64 n) {
    for (u64 t = u64(blockIdx.x) * blockDim.x + threadIdx.x; t < n; t += u64(gridDim.x) * blockDim.x) p[t] = t + 1;
}

// bit_vector_stat fast paths (config C3).  The generic kernel reads the whole 64-byte block from one thread
// (4 x 16-byte loads per lane: the slow L1 pattern of scripts/mb_gather.cu, 15 G queries/s).
// get: the one data word, 8 bytes per lane.
__global__ void __launch_bounds__(256) stat_get_kernel(RbImg v, const u64 *q, u64 n, unsigned char *out) {
    for (u64 t = u64(blockIdx.x) * blockDim.x + threadIdx.x; t < n; t += u64(gridDim.x) * blockDim.x) {
        const u64 x = q[t];
        bool bit = false;
        if (x < v.n_bits) {
            const u64 b = x / kRbBits;
            const u32 o = u32(x - b * kRbBits);
            u64 w;
            asm volatile("ld.global.nc.L1::no_allocate.L2::64B.u64 %0, [%1];" : "=l"(w) : "l"(v.blk + b * kRbWords + 1 + (o >> 6)));
            bit = (w >> (o & 63)) & 1ull;
        }
        out[t] = bit;
    }
}

// rank1: 2 lanes per query, each reads one 32-byte sector of the block with a 256-bit load
// (lane 0: header + words 0..2, lane 1: words 3..6), partial popcounts meet in one shuffle.
__global__ void __launch_bounds__(256) stat_rank_kernel(RbImg v, const u64 *q, u64 n, u64 *out) {
    const u32 half = threadIdx.x & 1;
    const u64 stride = (u64(gridDim.x) * blockDim.x) >> 1;
    for (u64 t0 = (u64(blockIdx.x) * blockDim.x + (threadIdx.x & ~31u)) >> 1; t0 < n; t0 += stride) {
        const u64 t = t0 + ((threadIdx.x & 31) >> 1);
        const bool valid = t < n && v.n_bits != 0;
        u64 part = 0;
        if (valid) {
            const u64 x = q[t];
            const u64 i = x < v.n_bits ? x : v.n_bits - 1;   // clamp (bit_vector.cpp:285)
            const u64 b = i / kRbBits;
            const u32 o = u32(i - b * kRbBits), wi = o >> 6;
            u32 r0, r1, r2, r3, r4, r5, r6, r7;
            asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v8.u32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7)
                         : "l"(v.blk + b * kRbWords + 4 * half));
            const u64 w[4] = {u64(r1) << 32 | r0, u64(r3) << 32 | r2, u64(r5) << 32 | r4, u64(r7) << 32 | r6};
            const u64 incl = (2ull << (o & 63)) - 1ull;       // bits 0..(o & 63)
#pragma unroll
            for (u32 k = 0; k < 4; ++k) {
                if (half == 0 && k == 0) { part = w[0]; continue; }   // the block's rank header
                const u32 word = 4 * half + k - 1;                     // data word index 0..6
                if (word < wi) part += __popcll(w[k]);
                else if (word == wi) part += __popcll(w[k] & incl);
            }
        }
        const u64 other = __shfl_xor_sync(0xffffffffu, part, 1);
        if (valid && half == 0) out[t] = part + other;
        if (t < n && v.n_bits == 0 && half == 0) out[t] = 0;
    }
}

// select1: 2 lanes per query.  Sample (block of every 2^s-th one; samples ~2 blocks apart) -> at most a couple of
// 8-byte header probes -> the block itself as two 32-byte sectors (256-bit loads), the in-block select split
// between the two lanes.  Out-of-range k (undefined in the reference, assert only) yields n_bits.
__global__ void __launch_bounds__(256) stat_select_kernel(RbImg v, const u64 *q, u64 n, u64 *out) {
    const u32 half = threadIdx.x & 1;
    const u64 stride = (u64(gridDim.x) * blockDim.x) >> 1;
    for (u64 t0 = (u64(blockIdx.x) * blockDim.x + (threadIdx.x & ~31u)) >> 1; t0 < n; t0 += stride) {
        const u64 t = t0 + ((threadIdx.x & 31) >> 1);
        const u64 x = t < n ? q[t] : 0;
        const bool ok = t < n && x >= 1 && x <= v.n_ones;
        u64 res = v.n_bits;
        u64 lo = 0, k0 = 0;
        if (ok) {
            k0 = x - 1;
            const u64 si = k0 >> v.s1;
            lo = __ldg(v.sel1 + si);
            u64 hi = __ldg(v.sel1 + si + 1);
            while (lo < hi) {   // last block in [lo, hi] with count-before <= k0
                const u64 mid = (lo + hi + 1) >> 1;
                u64 h;
                asm volatile("ld.global.nc.L1::no_allocate.L2::64B.u64 %0, [%1];" : "=l"(h) : "l"(v.blk + mid * kRbWords));
                if (h <= k0) lo = mid; else hi = mid - 1;
            }
        }
        u32 r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0, r5 = 0, r6 = 0, r7 = 0;
        if (ok)
            asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v8.u32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7)
                         : "l"(v.blk + lo * kRbWords + 4 * half));
        u64 w[4] = {u64(r1) << 32 | r0, u64(r3) << 32 | r2, u64(r5) << 32 | r4, u64(r7) << 32 | r6};
        const u64 header = __shfl_sync(0xffffffffu, w[0], threadIdx.x & 30);   // lane 0 of the pair holds it
        if (half == 0) w[0] = 0;                                              // not data
        const u32 mine = __popcll(w[0]) + __popcll(w[1]) + __popcll(w[2]) + __popcll(w[3]);
        const u32 first = __shfl_sync(0xffffffffu, mine, threadIdx.x & 30);   // ones in data words 0..2
        if (ok) {
            u32 k = u32(k0 - header);
            const bool owner = half == 0 ? k < first : k >= first;
            if (owner) {
                if (half) k -= first;
                const u64 base = lo * kRbBits + (half ? 3 * 64 : 0);
#pragma unroll
                for (u32 j = 0; j < 4; ++j) {
                    if (half == 0 && j == 0) continue;
                    const u32 c = __popcll(w[j]);
                    if (k < c) { res = base + (half ? j : j - 1) * 64 + select_in_word(w[j], k); break; }
                    k -= c;
                }
                out[t] = res;
            }
        } else if (t < n && half == 0) {
            out[t] = res;
        }
    }
}

// bit i set iff mix64(seed_mul + i) < threshold  (SURVEY.md 8d portable generator)
__global__ void synth_bits_kernel(u64 *words, u64 n_words, u64 n_bits, u64 threshold, u64 seed_mul) {
    u64 w = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    u64 x = 0;
    u64 base = w * 64;
    for (u32 b = 0; b < 64; ++b) {
        u64 i = base + b;
        if (i < n_bits && mix64(seed_mul + i) < threshold) x |= 1ull << b;
    }
    words[w] = x;
}

struct BitVec {
    int kind = 0, device = 0;
    u64 n_bits = 0, n_ones = 0;
    RbBuf stat;
    SdBuf sd;
    RrrBuf rrr;
    BvImg img{};
    cudaStream_t stream = nullptr;
    DeviceBuf<u64> d_q, d_out;
    std::mutex mu;
    mutable u32 aux_width = 64;   // get_int width of the next OP_GETINT launch (set under `mu` / by the device call)

    ~BitVec() { if (stream) cudaStreamDestroy(stream); }

    size_t bytes() const { return stat.bytes() + sd.bytes() + rrr.bytes(); }

    void build_from_device(int kind_, const u64 *d_plain, u64 n, int force_inverted) {
        kind = kind_;
        n_bits = n;
        CUDA_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        if (kind == GGQ_BV_STAT) { rb_build(&stat, d_plain, n, false, stream); n_ones = stat.img.n_ones; }
        else if (kind == GGQ_BV_SD) {
            sd_build(&sd, d_plain, n, force_inverted, stream);
            n_ones = sd.img.inverted ? n - sd.img.m : sd.img.m;
        } else if (kind == GGQ_BV_RRR) { rrr_build(&rrr, d_plain, n, stream); n_ones = rrr.img.n_ones; }
        else throw std::runtime_error("unknown bit vector kind");
        img.kind = kind; img.n_bits = n_bits; img.n_ones = n_ones;
        img.stat = stat.img; img.sd = sd.img; img.rrr = rrr.img;
    }

    template <int KIND>
    void launch_kind(int op, const u64 *q, u64 n, void *out, cudaStream_t st) const {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        u32 grid = u32(std::min<u64>((n + 255) / 256, u64(sms) * (KIND == GGQ_BV_RRR ? 6 : 8) * 4));
        if (!grid) return;
        if (op >= OP_RANK0) {
            u64 *o = static_cast<u64 *>(out);
            if (op == OP_RANK0) bv_aux_kernel<KIND, OP_RANK0><<<grid, 256, 0, st>>>(img, q, nullptr, 0, n, o);
            else if (op == OP_SELECT0) bv_aux_kernel<KIND, OP_SELECT0><<<grid, 256, 0, st>>>(img, q, nullptr, 0, n, o);
            else if (op == OP_GETINT) bv_aux_kernel<KIND, OP_GETINT><<<grid, 256, 0, st>>>(img, q, nullptr, aux_width, n, o);
            else if (op == OP_MAPRANK) bv_aux_kernel<KIND, OP_MAPRANK><<<grid, 256, 0, st>>>(img, q, nullptr, 0, n, o);
            else bv_aux_kernel<KIND, OP_WORDS><<<grid, 256, 0, st>>>(img, nullptr, nullptr, 0, n, o);
            CUDA_CHECK(cudaGetLastError());
            return;
        }
        if (KIND == GGQ_BV_STAT && op == OP_GET) {
            stat_get_kernel<<<grid, 256, 0, st>>>(img.stat, q, n, static_cast<unsigned char *>(out));
        } else if (KIND == GGQ_BV_STAT && op == OP_RANK1) {
            u32 g2 = u32(std::min<u64>((2 * n + 255) / 256, u64(sms) * 32));
            stat_rank_kernel<<<g2, 256, 0, st>>>(img.stat, q, n, static_cast<u64 *>(out));
        } else if (KIND == GGQ_BV_STAT && op == OP_SELECT1) {
            u32 g2 = u32(std::min<u64>((2 * n + 255) / 256, u64(sms) * 32));
            stat_select_kernel<<<g2, 256, 0, st>>>(img.stat, q, n, static_cast<u64 *>(out));
        } else
        if (op == OP_RANK1) bv_query_kernel<KIND, OP_RANK1, u64><<<grid, 256, 0, st>>>(img, q, n, static_cast<u64 *>(out));
        else if (op == OP_SELECT1) bv_query_kernel<KIND, OP_SELECT1, u64><<<grid, 256, 0, st>>>(img, q, n, static_cast<u64 *>(out));
        else bv_query_kernel<KIND, OP_GET, unsigned char><<<grid, 256, 0, st>>>(img, q, n, static_cast<unsigned char *>(out));
        CUDA_CHECK(cudaGetLastError());
    }

    void launch(int op, const u64 *q, u64 n, void *out, cudaStream_t st) const {
        if (kind == GGQ_BV_STAT) launch_kind<GGQ_BV_STAT>(op, q, n, out, st);
        else if (kind == GGQ_BV_SD) launch_kind<GGQ_BV_SD>(op, q, n, out, st);
        else launch_kind<GGQ_BV_RRR>(op, q, n, out, st);
    }

    // positions of the set bits, ascending (call_ones / the sparse branch of bit_vector_*::call_ones): select1(1..n_ones)
    void ones_host(u64 *out) {
        if (!n_ones) return;
        std::lock_guard<std::mutex> lk(mu);
        CUDA_CHECK(cudaSetDevice(device));
        if (d_q.n < n_ones) { d_q.alloc(n_ones); d_out.alloc(n_ones); }
        iota1_kernel<<<u32(std::min<u64>((n_ones + 255) / 256, 4096)), 256, 0, stream>>>(d_q.p, n_ones);
        launch(OP_SELECT1, d_q.p, n_ones, d_out.p, stream);
        CUDA_CHECK(cudaMemcpyAsync(out, d_out.p, n_ones * 8, cudaMemcpyDeviceToHost, stream));
        CUDA_CHECK(cudaStreamSynchronize(stream));
    }

    void query_host(int op, const u64 *q, u64 n, void *out, u32 width = 64) {
        if (!n) return;
        std::lock_guard<std::mutex> lk(mu);
        aux_width = width;
        CUDA_CHECK(cudaSetDevice(device));
        if (d_q.n < n) { d_q.alloc(n); d_out.alloc(n); }
        if (q) CUDA_CHECK(cudaMemcpyAsync(d_q.p, q, n * 8, cudaMemcpyHostToDevice, stream));
        launch(op, d_q.p, n, d_out.p, stream);
        CUDA_CHECK(cudaMemcpyAsync(out, d_out.p, op == OP_GET ? n : n * 8, cudaMemcpyDeviceToHost, stream));
        CUDA_CHECK(cudaStreamSynchronize(stream));
    }
};

inline BitVec *bv_from_host_words(int kind, const u64 *words, u64 n_bits, int device, int force_inverted) {
    CUDA_CHECK(cudaSetDevice(device));
    std::unique_ptr<BitVec> v(new BitVec());
    v->device = device;
    u64 n_words = (n_bits + 63) / 64;
    DeviceBuf<u64> plain;
    plain.alloc(std::max<u64>(n_words, 1));
    if (n_words) {
        CUDA_CHECK(cudaMemcpy(plain.p, words, n_words * 8, cudaMemcpyHostToDevice));
        if (n_bits & 63) {  // clear padding
            u64 last = words[n_words - 1] & ((1ull << (n_bits & 63)) - 1ull);
            CUDA_CHECK(cudaMemcpy(plain.p + n_words - 1, &last, 8, cudaMemcpyHostToDevice));
        }
    }
    v->build_from_device(kind, plain.p, n_bits, force_inverted);
    return v.release();
}

}  // namespace ggq

struct ggq_bv : ggq::BitVec {};

extern "C" {

int ggq_bv_build(int kind, const uint64_t *words, uint64_t n_bits, int device, ggq_bv **out) {
    GGQ_TRY
    if (!out || (!words && n_bits) || kind < 0 || kind > 2) return fail(GGQ_ERR_ARG, "bad argument");
    *out = static_cast<ggq_bv *>(bv_from_host_words(kind, reinterpret_cast<const u64 *>(words), n_bits, device, -1));
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_bv_load_buffer(int kind, const uint8_t *buf, size_t len, size_t *consumed, int device, ggq_bv **out) {
    GGQ_TRY
    if (!out || (!buf && len) || kind < 0 || kind > 2) return fail(GGQ_ERR_ARG, "bad argument");
    Reader in(buf, len);
    PlainBits bits;
    int force_inv = -1;
    if (kind == GGQ_BV_STAT) load_stat(in, &bits);
    else if (kind == GGQ_BV_RRR) load_rrr63(in, &bits);
    else {
        SdColumn c;
        load_sd(in, &c);
        sd_to_plain(c, &bits);
        force_inv = c.inverted ? 1 : 0;   // keep the file's representation choice (is_inverted())
    }
    if (consumed) *consumed = in.consumed();
    std::vector<u64> w = plain_to_words64(bits);
    *out = static_cast<ggq_bv *>(bv_from_host_words(kind, w.data(), bits.size, device, force_inv));
    return GGQ_OK;
    GGQ_CATCH
}

int ggq_bv_synthesize(int kind, uint64_t n_bits, uint64_t threshold, uint64_t seed, int device, ggq_bv **out) {
    GGQ_TRY
    if (!out || kind < 0 || kind > 2) return fail(GGQ_ERR_ARG, "bad argument");
    CUDA_CHECK(cudaSetDevice(device));
    std::unique_ptr<BitVec> v(new BitVec());
    v->device = device;
    u64 n_words = (n_bits + 63) / 64;
    DeviceBuf<u64> plain;
    plain.alloc(std::max<u64>(n_words, 1));
    if (n_words) {
        synth_bits_kernel<<<u32((n_words + 255) / 256), 256>>>(plain.p, n_words, n_bits, threshold, seed * 0x9E3779B97F4A7C15ULL);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaDeviceSynchronize());
    }
    v->build_from_device(kind, plain.p, n_bits, -1);
    *out = static_cast<ggq_bv *>(v.release());
    return GGQ_OK;
    GGQ_CATCH
}

void ggq_bv_free(ggq_bv *v) {
    if (!v) return;
    cudaSetDevice(v->device);
    delete static_cast<ggq::BitVec *>(v);
}
uint64_t ggq_bv_size(const ggq_bv *v) { return v->n_bits; }
uint64_t ggq_bv_num_set_bits(const ggq_bv *v) { return v->n_ones; }
int ggq_bv_is_inverted(const ggq_bv *v) { return v->kind == GGQ_BV_SD ? int(v->sd.img.inverted) : 0; }
uint64_t ggq_bv_device_bytes(const ggq_bv *v) { return v->bytes(); }

static int bv_host(const ggq_bv *v, int op, const uint64_t *q, uint64_t n, void *out) {
    GGQ_TRY
    if (!v || (!q && n) || (!out && n)) return fail(GGQ_ERR_ARG, "bad argument");
    const_cast<ggq_bv *>(v)->query_host(op, reinterpret_cast<const u64 *>(q), n, out);
    return GGQ_OK;
    GGQ_CATCH
}
int ggq_bv_rank1(const ggq_bv *v, const uint64_t *q, uint64_t n, uint64_t *out) { return bv_host(v, OP_RANK1, q, n, out); }
int ggq_bv_select1(const ggq_bv *v, const uint64_t *q, uint64_t n, uint64_t *out) { return bv_host(v, OP_SELECT1, q, n, out); }
int ggq_bv_get(const ggq_bv *v, const uint64_t *q, uint64_t n, uint8_t *out) { return bv_host(v, OP_GET, q, n, out); }
int ggq_bv_rank0(const ggq_bv *v, const uint64_t *q, uint64_t n, uint64_t *out) { return bv_host(v, OP_RANK0, q, n, out); }
int ggq_bv_select0(const ggq_bv *v, const uint64_t *q, uint64_t n, uint64_t *out) { return bv_host(v, OP_SELECT0, q, n, out); }
int ggq_bv_get_int(const ggq_bv *v, const uint64_t *q, uint64_t n, uint32_t width, uint64_t *out) {
    GGQ_TRY
    if (!v || (!q && n) || (!out && n) || width < 1 || width > 64) return fail(GGQ_ERR_ARG, "bad argument");
    const_cast<ggq_bv *>(v)->query_host(OP_GETINT, reinterpret_cast<const u64 *>(q), n, out, width);
    return GGQ_OK;
    GGQ_CATCH
}
int ggq_bv_to_words(const ggq_bv *v, uint64_t *words) {
    GGQ_TRY
    if (!v || (!words && v->n_bits)) return fail(GGQ_ERR_ARG, "bad argument");
    const_cast<ggq_bv *>(v)->query_host(OP_WORDS, nullptr, (v->n_bits + 63) / 64, words);
    return GGQ_OK;
    GGQ_CATCH
}
int ggq_bv_call_ones_into(const ggq_bv *v, uint64_t *positions, uint64_t capacity, uint64_t *n_ones) {
    GGQ_TRY
    if (!v || !n_ones) return fail(GGQ_ERR_ARG, "bad argument");
    *n_ones = v->n_ones;
    if (capacity < v->n_ones || (!positions && v->n_ones)) return fail(GGQ_ERR_ARG, "capacity too small");
    const_cast<ggq_bv *>(v)->ones_host(reinterpret_cast<u64 *>(positions));
    return GGQ_OK;
    GGQ_CATCH
}

static int bv_dev(const ggq_bv *v, int op, const uint64_t *d_q, uint64_t n, void *d_out, void *stream) {
    GGQ_TRY
    if (!v || (!d_q && n) || (!d_out && n)) return fail(GGQ_ERR_ARG, "bad argument");
    CUDA_CHECK(cudaSetDevice(v->device));
    v->launch(op, reinterpret_cast<const u64 *>(d_q), n, d_out, static_cast<cudaStream_t>(stream));
    return GGQ_OK;
    GGQ_CATCH
}
int ggq_bv_rank1_device(const ggq_bv *v, const uint64_t *d_q, uint64_t n, uint64_t *d_out, void *stream) { return bv_dev(v, OP_RANK1, d_q, n, d_out, stream); }
int ggq_bv_select1_device(const ggq_bv *v, const uint64_t *d_q, uint64_t n, uint64_t *d_out, void *stream) { return bv_dev(v, OP_SELECT1, d_q, n, d_out, stream); }
int ggq_bv_get_device(const ggq_bv *v, const uint64_t *d_q, uint64_t n, uint8_t *d_out, void *stream) { return bv_dev(v, OP_GET, d_q, n, d_out, stream); }
int ggq_map_rank1(const ggq_bv *mask, const uint64_t *graph_idx, uint64_t n, uint64_t *anno_idx, uint64_t *n_unmapped) {
    int rc = bv_host(mask, OP_MAPRANK, graph_idx, n, anno_idx);
    if (rc == GGQ_OK && n_unmapped) {
        uint64_t bad = 0;
        for (uint64_t i = 0; i < n; ++i) bad += anno_idx[i] == ~uint64_t(0);
        *n_unmapped = bad;
    }
    return rc;
}
int ggq_map_rank1_device(const ggq_bv *mask, const uint64_t *d_graph_idx, uint64_t n, uint64_t *d_anno_idx, void *stream) {
    return bv_dev(mask, OP_MAPRANK, d_graph_idx, n, d_anno_idx, stream);
}
int ggq_bv_rank0_device(const ggq_bv *v, const uint64_t *d_q, uint64_t n, uint64_t *d_out, void *stream) { return bv_dev(v, OP_RANK0, d_q, n, d_out, stream); }
int ggq_bv_select0_device(const ggq_bv *v, const uint64_t *d_q, uint64_t n, uint64_t *d_out, void *stream) { return bv_dev(v, OP_SELECT0, d_q, n, d_out, stream); }

}  // extern "C"
