// Elias-Fano device layout for bit_vector_sd (common/bit_vector.cpp:572-590, 690-718; sdsl sd_vector.hpp).
// Stored vector S = the logical bits, or their complement when `inverted` (more than half ones).
// The m ones of S at positions p_0 < ... < p_{m-1}:  low[k] = p_k & (2^wl - 1) packed wl bits each;
// high has a one at (p_k >> wl) + k, as RankedBits with select1 and select0 samples.
#pragma once

namespace ggq {

struct SdImg {
    RbImg high;
    const u64 *low;   // packed, + 1 pad word
    u64 n;            // logical size
    u64 m;            // ones of the stored vector
    u32 wl;
    u32 inverted;
};

__device__ __forceinline__ u64 sd_low(const SdImg &v, u64 k) {
    if (v.wl == 0) return 0;
    u64 bp = k * v.wl;
    u64 w = bp >> 6;
    u32 o = u32(bp & 63);
    u64 x = __ldg(v.low + w) >> o;
    if (o + v.wl > 64) x |= __ldg(v.low + w + 1) << (64 - o);
    return x & ((1ull << v.wl) - 1ull);
}

// data word `widx` (positions widx*64 .. +63) of an rb vector; 448 = 7 * 64 so words never straddle blocks
__device__ __forceinline__ u64 rb_data_word(const RbImg &v, u64 widx) {
    u64 b = widx / 7;
    return __ldg(v.blk + b * kRbWords + 1 + (widx - b * 7));
}

// position of the k-th (0-based) one of the stored vector
__device__ __forceinline__ u64 sd_stored_select(const SdImg &v, u64 k) {
    u64 h = rb_select<false>(v.high, k);
    return ((h - k) << v.wl) | sd_low(v, k);
}

// ones of the stored vector in [0, i] (i < n); *hit = stored bit i
__device__ __forceinline__ u64 sd_stored_rank_incl(const SdImg &v, u64 i, bool *hit) {
    const u64 hb = i >> v.wl;
    const u64 lo = v.wl ? (i & ((1ull << v.wl) - 1ull)) : 0ull;
    u64 pos = 0, idx = 0;
    if (hb) {
        u64 z = rb_select<true>(v.high, hb - 1);
        pos = z + 1;
        idx = pos - hb;
    }
    bool eq = false;
    u64 widx = pos >> 6;
    u32 avail = 64 - u32(pos & 63);
    u64 w = rb_data_word(v.high, widx) >> (pos & 63);
    for (;;) {
        if (!avail) { ++widx; w = rb_data_word(v.high, widx); avail = 64; }
        if (!(w & 1ull)) break;
        u64 l = sd_low(v, idx);
        if (l > lo) break;
        eq = (l == lo);
        ++idx;
        w >>= 1;
        --avail;
    }
    *hit = eq;
    return idx;
}

__device__ __forceinline__ bool sd_get(const SdImg &v, u64 i) {
    if (i >= v.n) return false;
    bool hit;
    sd_stored_rank_incl(v, i, &hit);
    return hit != bool(v.inverted);
}

__device__ __forceinline__ u64 sd_rank1(const SdImg &v, u64 i) {
    if (v.n == 0) return 0;
    if (i >= v.n) i = v.n - 1;
    bool hit;
    u64 r = sd_stored_rank_incl(v, i, &hit);
    return v.inverted ? (i + 1) - r : r;
}

// k in [1, logical ones]
__device__ __forceinline__ u64 sd_select1(const SdImg &v, u64 k) {
    if (!v.inverted) return sd_stored_select(v, k - 1);
    // k-th zero of the stored vector: smallest j in [0, m] with (zeros before stored one #j) >= k
    u64 lo = 0, hi = v.m;
    while (lo < hi) {
        u64 mid = (lo + hi) >> 1;
        u64 zeros_before = sd_stored_select(v, mid) - mid;
        if (zeros_before >= k) hi = mid; else lo = mid + 1;
    }
    return k - 1 + lo;
}

// ---------------------------------------------------------------------------------------------
// construction from an rb over the stored bits: thread per rb block enumerates its ones
// ---------------------------------------------------------------------------------------------
__global__ void sd_scatter_kernel(RbImg src, u32 wl, u64 *low, u64 *high_plain) {
    u64 b = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= src.n_blk) return;
    const u64 *blk = src.blk + b * kRbWords;
    u64 idx = blk[0];
    const u64 mask = wl ? ((1ull << wl) - 1ull) : 0ull;
    for (u32 j = 0; j < 7; ++j) {
        u64 x = blk[1 + j];
        while (x) {
            u32 t = __ffsll((long long)x) - 1;
            x &= x - 1;
            u64 p = b * kRbBits + j * 64 + t;
            u64 hp = (p >> wl) + idx;
            atomicOr(high_plain + (hp >> 6), 1ull << (hp & 63));
            if (wl) {
                u64 bp = idx * wl;
                u64 l = p & mask;
                u32 o = u32(bp & 63);
                atomicOr(low + (bp >> 6), l << o);
                if (o + wl > 64) atomicOr(low + (bp >> 6) + 1, l >> (64 - o));
            }
            ++idx;
        }
    }
}

__global__ void complement_kernel(const u64 *in, u64 *out, u64 n_words, u64 n_bits) {
    u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n_words) return;
    u64 x = ~in[i];
    if (i == n_words - 1 && (n_bits & 63)) x &= (1ull << (n_bits & 63)) - 1ull;
    out[i] = x;
}

__global__ void popcount_kernel(const u64 *in, u64 n_words, u64 *total) {
    u64 i = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    u32 c = i < n_words ? __popcll(in[i]) : 0;
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(total, u64(c));
}

struct SdBuf {
    RbBuf high;
    DeviceBuf<u64> low;
    SdImg img{};
    size_t bytes() const { return high.bytes() + low.bytes(); }
};

inline u64 device_popcount(const u64 *d_words, u64 n_words, cudaStream_t st = nullptr) {
    if (!n_words) return 0;
    DeviceBuf<u64> tot;
    tot.alloc(1);
    CUDA_CHECK(cudaMemsetAsync(tot.p, 0, 8, st));
    popcount_kernel<<<u32((n_words + 255) / 256), 256, 0, st>>>(d_words, n_words, tot.p);
    u64 h = 0;
    CUDA_CHECK(cudaMemcpyAsync(&h, tot.p, 8, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    return h;
}

// force_inverted: -1 = the reference's rule (popcount > size / 2, bit_vector.cpp:576), else the given flag
inline void sd_build(SdBuf *out, const u64 *d_plain, u64 n_bits, int force_inverted, cudaStream_t st = nullptr) {
    u64 n_words = (n_bits + 63) / 64;
    u64 ones = device_popcount(d_plain, n_words, st);
    bool inv = force_inverted < 0 ? (ones > n_bits / 2) : (force_inverted != 0);
    DeviceBuf<u64> comp;
    const u64 *stored = d_plain;
    if (inv) {
        comp.alloc(std::max<u64>(n_words, 1));
        if (n_words) complement_kernel<<<u32((n_words + 255) / 256), 256, 0, st>>>(d_plain, comp.p, n_words, n_bits);
        stored = comp.p;
    }
    u64 m = inv ? n_bits - ones : ones;
    u32 wl = 0;
    {
        u64 ratio = n_bits / std::max<u64>(m, 1);
        while (wl < 62 && (2ull << wl) <= ratio) ++wl;
    }
    RbBuf src;
    rb_build(&src, stored, n_bits, false, st);
    comp.release();
    u64 high_bits = m + (n_bits >> wl) + 1;
    u64 high_words = (high_bits + 63) / 64;
    DeviceBuf<u64> high_plain;
    high_plain.alloc(high_words);
    CUDA_CHECK(cudaMemsetAsync(high_plain.p, 0, high_words * 8, st));
    u64 low_words = (m * wl + 63) / 64 + 1;
    out->low.alloc(low_words);
    CUDA_CHECK(cudaMemsetAsync(out->low.p, 0, low_words * 8, st));
    sd_scatter_kernel<<<u32((src.img.n_blk + 127) / 128), 128, 0, st>>>(src.img, wl, out->low.p, high_plain.p);
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaStreamSynchronize(st));
    src = RbBuf();
    rb_build(&out->high, high_plain.p, high_bits, true, st);
    out->img.high = out->high.img;
    out->img.low = out->low.p;
    out->img.n = n_bits;
    out->img.m = m;
    out->img.wl = wl;
    out->img.inverted = inv;
}

}  // namespace ggq
