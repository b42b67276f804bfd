// K5: GPU-resident bit_vector_{stat, sd, rrr<63>} (common/bit_vector.hpp:12-225) — device layouts with
// 64-bit aligned, INTERLEAVED rank directories and sampled select.
//
//   RankedBits ("rb", used by stat and by the Elias-Fano high part):
//       64-byte blocks = { u64 ones_before_block, u64 data[7] }  (448 positions, 12.5 % overhead);
//       bit test + rank = ONE 64-byte transaction.  select1 / select0: one u32 sample per 2^s ones
//       (zeros), s chosen per vector so that consecutive samples are ~2 blocks apart, then a
//       <= 3-step binary search over block headers and an in-block __popcll / __fns scan.
//   stat : rb over the plain bits            (reference: sdsl rank_support_v5 + select_support_mcl,
//                                             bit_vector.cpp:281-304, rank_support_v5.hpp:116-134)
//   sd   : Elias-Fano of the ones — or of the zeros when popcount > size/2, exactly the reference's
//          inversion rule (bit_vector.cpp:572-590): low parts packed wl bits each, high part as rb
//          (reference: sdsl sd_vector.hpp:192-227,487-582)
//   rrr  : 63-bit blocks, 16 per superblock; a 32-byte superblock header { u64 ones_before,
//          u64 offset_bit_pointer, 16 x 6-bit classes } + a packed stream of combinatorial offsets;
//          access/rank = header sector + offset word(s), block decoded with a binomial walk from a
//          64x64 table staged in shared memory (reference: sdsl rrr_vector.hpp:257-277,445-590).
//
// Semantics (bit-exact with the reference): rank1(i) = ones in [0, i], i >= size clamps
// (bit_vector.cpp:285,692,860); select1(k) = position of the k-th one, k in [1, ones]; get(i) = bit i.
#pragma once
#include <cub/device/device_scan.cuh>

#include "ggq_engine.hpp"

namespace ggq {

constexpr u32 kRbBits = 448;          // positions per 64-byte block
constexpr u32 kRbWords = 8;           // u64 words per block (header + 7 data)
constexpr u32 kRbSpanBlocks = 2;      // target distance (blocks) between consecutive select samples

struct RbImg {
    const u64 *blk;     // n_blk * 8 words
    const u32 *sel1;    // block of every (j << s1)-th one, + sentinel
    const u32 *sel0;    // block of every (j << s0)-th zero, + sentinel (may be null)
    u64 n_bits, n_ones, n_blk;
    u32 s1, s0;
};

__device__ __forceinline__ void rb_load(const RbImg &v, u64 b, u64 (&w)[8]) {
    const ulonglong2 *p = reinterpret_cast<const ulonglong2 *>(v.blk + b * kRbWords);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        ulonglong2 t;
        asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(t.x), "=l"(t.y) : "l"(p + i));
        w[2 * i] = t.x;
        w[2 * i + 1] = t.y;
    }
}

__device__ __forceinline__ u64 rb_header(const RbImg &v, u64 b) { return __ldg(v.blk + b * kRbWords); }

// ones in [0, i) and the bit at i; i < n_bits
__device__ __forceinline__ u64 rb_rank_excl_bit(const RbImg &v, u64 i, bool *bit) {
    u64 b = i / kRbBits;
    u32 o = u32(i - b * kRbBits);
    u64 w[8];
    rb_load(v, b, w);
    u64 r = w[0];
    u32 wi = o >> 6;
#pragma unroll
    for (u32 k = 0; k < 7; ++k)
        if (k < wi) r += __popcll(w[1 + k]);
    u64 x = 0;
#pragma unroll
    for (u32 k = 0; k < 7; ++k)
        if (k == wi) x = w[1 + k];
    *bit = (x >> (o & 63)) & 1ull;
    return r + __popcll(x & ((1ull << (o & 63)) - 1ull));
}

// position (0-based) of the k-th (0-based) set bit of x; k < popcll(x)
__device__ __forceinline__ u32 select_in_word(u64 x, u32 k) {
    u32 lo = u32(x), c = __popc(lo);
    if (k < c) return __fns(lo, 0, int(k) + 1);
    return 32 + __fns(u32(x >> 32), 0, int(k - c) + 1);
}

// position of the q-th (0-based) one (ZERO = false) or zero (ZERO = true); q < n_ones (n_zeros)
template <bool ZERO>
__device__ __forceinline__ u64 rb_select(const RbImg &v, u64 q) {
    const u32 *sel = ZERO ? v.sel0 : v.sel1;
    const u32 s = ZERO ? v.s0 : v.s1;
    u64 si = q >> s;
    u64 lo = __ldg(sel + si), hi = __ldg(sel + si + 1);
    // last block in [lo, hi] with count-before <= q
    while (lo < hi) {
        u64 mid = (lo + hi + 1) >> 1;
        u64 h = rb_header(v, mid);
        u64 before = ZERO ? mid * kRbBits - h : h;
        if (before <= q) lo = mid; else hi = mid - 1;
    }
    u64 w[8];
    rb_load(v, lo, w);
    u64 before = ZERO ? lo * kRbBits - w[0] : w[0];
    u32 k = u32(q - before);
    u64 base = lo * kRbBits;
#pragma unroll
    for (u32 j = 0; j < 7; ++j) {
        u64 x = ZERO ? ~w[1 + j] : w[1 + j];
        u32 c = __popcll(x);
        if (k < c) return base + j * 64 + select_in_word(x, k);
        k -= c;
    }
    return v.n_bits;  // unreachable for valid q
}

// ---------------------------------------------------------------------------------------------
// rb construction from plain device words (LSB-first u64, padding bits zero)
// ---------------------------------------------------------------------------------------------
__global__ void rb_count_kernel(const u64 *plain, u64 n_words, u64 n_blk, u64 *cnt) {
    u64 b = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= n_blk) return;
    u32 c = 0;
    for (u32 k = 0; k < 7; ++k) {
        u64 wi = b * 7 + k;
        if (wi < n_words) c += __popcll(plain[wi]);
    }
    cnt[b] = c;
}

__global__ void rb_fill_kernel(const u64 *plain, u64 n_words, u64 n_blk, const u64 *before, u64 *blk) {
    u64 b = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= n_blk) return;
    u64 *dst = blk + b * kRbWords;
    dst[0] = before[b];
    for (u32 k = 0; k < 7; ++k) {
        u64 wi = b * 7 + k;
        dst[1 + k] = wi < n_words ? plain[wi] : 0ull;
    }
}

// sel[j] = block that contains the (j << s)-th one (zero); sentinel sel[n_samples] = last block
template <bool ZERO>
__global__ void rb_sample_kernel(const u64 *blk, u64 n_blk, u64 n_bits, u32 s, u32 *sel, u64 n_samples, u64 total) {
    u64 b = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= n_blk) return;
    u64 h0 = blk[b * kRbWords];
    u64 lo, hi;
    if (!ZERO) {
        lo = h0;
        u32 c = 0;
        for (u32 k = 0; k < 7; ++k) c += __popcll(blk[b * kRbWords + 1 + k]);
        hi = h0 + c;
    } else {
        u64 start = b * kRbBits, end = min(n_bits, start + kRbBits);
        if (start >= end) { lo = hi = 0; }
        else {
            u32 c = 0;
            for (u32 k = 0; k < 7; ++k) c += __popcll(blk[b * kRbWords + 1 + k]);
            lo = start - h0;
            hi = end - (h0 + c);
        }
    }
    // multiples of 2^s in [lo, hi)
    u64 first = (lo + (1ull << s) - 1) >> s;
    for (u64 j = first; (j << s) < hi; ++j) sel[j] = u32(b);
    if (b == n_blk - 1)  // tail: samples at or beyond `total` (incl. the sentinel) point at the last block
        for (u64 j = (total + (1ull << s) - 1) >> s; j <= n_samples; ++j) sel[j] = u32(b);
}

struct RbBuf {
    DeviceBuf<u64> blk;
    DeviceBuf<u32> sel1, sel0;
    RbImg img{};
    size_t bytes() const { return blk.bytes() + sel1.bytes() + sel0.bytes(); }
};

inline u32 pick_sample_shift(u64 count, u64 n_blk) {
    // consecutive samples ~kRbSpanBlocks blocks apart on average: 2^s ~ count * span / n_blk
    double per = n_blk ? double(count) * kRbSpanBlocks / double(n_blk) : 1.0;
    u32 s = 0;
    while (s < 16 && double(2ull << s) <= per) ++s;
    return s;
}

// d_plain: n_words = ceil(n_bits / 64) words on the device
inline void rb_build(RbBuf *out, const u64 *d_plain, u64 n_bits, bool want_sel0, cudaStream_t st = nullptr) {
    u64 n_words = (n_bits + 63) / 64;
    u64 n_blk = n_bits / kRbBits + 1;   // + sentinel so that rank(n_bits) and clamped queries read a valid block
    if (n_blk > 0xFFFFFFF0ull) throw LimitError("bit vector too long");
    DeviceBuf<u64> cnt, before;
    cnt.alloc(n_blk + 1);
    before.alloc(n_blk + 1);
    u32 g = u32((n_blk + 255) / 256);
    rb_count_kernel<<<g, 256, 0, st>>>(d_plain, n_words, n_blk, cnt.p);
    CUDA_CHECK(cudaMemsetAsync(cnt.p + n_blk, 0, 8, st));
    size_t need = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, need, cnt.p, before.p, n_blk + 1, st);
    DeviceBuf<unsigned char> tmp;
    tmp.alloc(need + 16);
    cub::DeviceScan::ExclusiveSum(tmp.p, need, cnt.p, before.p, n_blk + 1, st);
    out->blk.alloc(n_blk * kRbWords);
    rb_fill_kernel<<<g, 256, 0, st>>>(d_plain, n_words, n_blk, before.p, out->blk.p);
    CUDA_CHECK(cudaGetLastError());
    u64 n_ones = 0;
    CUDA_CHECK(cudaMemcpyAsync(&n_ones, before.p + n_blk, 8, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    RbImg &im = out->img;
    im.blk = out->blk.p;
    im.n_bits = n_bits; im.n_ones = n_ones; im.n_blk = n_blk;
    im.s1 = pick_sample_shift(n_ones, n_blk);
    u64 ns1 = (n_ones >> im.s1) + 1;
    out->sel1.alloc(ns1 + 1);
    CUDA_CHECK(cudaMemsetAsync(out->sel1.p, 0, (ns1 + 1) * 4, st));
    rb_sample_kernel<false><<<g, 256, 0, st>>>(out->blk.p, n_blk, n_bits, im.s1, out->sel1.p, ns1, n_ones);
    im.sel1 = out->sel1.p;
    im.sel0 = nullptr; im.s0 = 0;
    if (want_sel0) {
        u64 n_zeros = n_bits - n_ones;
        im.s0 = pick_sample_shift(n_zeros, n_blk);
        u64 ns0 = (n_zeros >> im.s0) + 1;
        out->sel0.alloc(ns0 + 1);
        CUDA_CHECK(cudaMemsetAsync(out->sel0.p, 0, (ns0 + 1) * 4, st));
        rb_sample_kernel<true><<<g, 256, 0, st>>>(out->blk.p, n_blk, n_bits, im.s0, out->sel0.p, ns0, n_zeros);
        im.sel0 = out->sel0.p;
    }
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaStreamSynchronize(st));
}

// PlainBits (u32 words, host parsers) -> LSB-first u64 words
inline std::vector<u64> plain_to_words64(const PlainBits &b) {
    u64 n_words = (b.size + 63) / 64;
    std::vector<u64> w(std::max<u64>(n_words, 1), 0);
    for (u64 i = 0; i < n_words; ++i) {
        u64 lo = b.w[2 * i], hi = (2 * i + 1 < b.w.size()) ? b.w[2 * i + 1] : 0;
        w[i] = lo | (hi << 32);
    }
    return w;
}

}  // namespace ggq

#include "ggq_bv_sd.cuh"
#include "ggq_bv_rrr.cuh"
