// RRR device layout for bit_vector_rrr<63> (common/bit_vector.cpp:806-983; sdsl rrr_vector.hpp:72-680).
// 63-bit blocks (the reference's t_bs), 16 blocks per superblock.  Per superblock ONE 32-byte sector:
//     { u64 ones_before, u64 offset_bit_pointer, u8 class[16] }        class = popcount of the block
// followed (in a separate packed stream) by the blocks' combinatorial offsets, space[class] bits each,
// space[k] = ceil(log2 C(63, k)).  A block with ones at c_1 < ... < c_k has offset N = sum_t C(c_t, t)
// (combinatorial number system); decoding walks i = 62 .. 0 and takes bit i iff C(i, t) <= N.
// access / rank = header sector + 1-2 offset words; select = sample + binary search over headers.
#pragma once

namespace ggq {

constexpr u32 kRrrBlock = 63;
constexpr u32 kRrrPerSb = 16;
constexpr u32 kRrrSbBits = kRrrBlock * kRrrPerSb;   // 1008
constexpr u32 kRrrSpanSb = 8;

struct RrrTables {
    u64 binom[64][64];
    unsigned char space[64];
};
__device__ RrrTables g_rrr_tables;

inline void rrr_init_tables() {
    static bool done[64] = {};
    int dev = 0;
    CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 64 && done[dev]) return;
    static RrrTables h;
    for (int n = 0; n < 64; ++n)
        for (int k = 0; k < 64; ++k) h.binom[n][k] = (k == 0) ? 1 : 0;
    for (int n = 1; n < 64; ++n)
        for (int k = 1; k <= n; ++k) h.binom[n][k] = h.binom[n - 1][k - 1] + h.binom[n - 1][k];
    for (int k = 0; k < 64; ++k) {
        u64 c = h.binom[63][k];
        h.space[k] = c <= 1 ? 0 : (unsigned char)(64 - __builtin_clzll(c - 1));
    }
    CUDA_CHECK(cudaMemcpyToSymbol(g_rrr_tables, &h, sizeof(h)));
    if (dev < 64) done[dev] = true;
}

struct RrrImg {
    const ulonglong4 *hdr;   // n_sb headers of 32 bytes
    const u64 *offs;         // packed offsets (+ 1 pad word)
    const u32 *sel1;         // superblock of every (j << s1)-th one, + sentinel
    u64 n_bits, n_ones, n_sb;
    u32 s1;
};

__device__ __forceinline__ u64 read_bits(const u64 *a, u64 bp, u32 len) {
    if (!len) return 0;
    u64 w = bp >> 6;
    u32 o = u32(bp & 63);
    u64 x = __ldg(a + w) >> o;
    if (o + len > 64) x |= __ldg(a + w + 1) << (64 - o);
    return len == 64 ? x : x & ((1ull << len) - 1ull);
}

__device__ __forceinline__ u32 hdr_class(const ulonglong4 &h, u32 j) {
    u64 w = j < 8 ? h.z : h.w;
    return u32(w >> ((j & 7) * 8)) & 0xffu;
}

// Decode walk.  Returns the number of ones at positions > o; *bit = bit o.
__device__ __forceinline__ u32 rrr_ones_above(const RrrTables &T, u32 k, u64 N, u32 o, bool *bit) {
    if (k == 0) { *bit = false; return 0; }
    if (k == kRrrBlock) { *bit = true; return kRrrBlock - 1 - o; }
    u32 t = k;
    for (u32 i = kRrrBlock - 1; i > o; --i) {
        u64 c = T.binom[i][t];
        if (t && c <= N) { N -= c; --t; }
    }
    *bit = t && T.binom[o][t] <= N;
    return k - t;
}

// position (0..62) of the r-th (0-based, from the low end) one of a block of class k with offset N; r < k
__device__ __forceinline__ u32 rrr_select_in_block(const RrrTables &T, u32 k, u64 N, u32 r) {
    if (k == kRrrBlock) return r;
    u32 t = k;
    const u32 want = r + 1;   // stop when the one about to be taken is the (r+1)-th from the bottom, i.e. t == want
    for (u32 i = kRrrBlock; i-- > 0;) {
        u64 c = T.binom[i][t];
        if (c <= N) {
            if (t == want) return i;
            N -= c;
            --t;
        }
    }
    return 0;  // unreachable
}

__device__ __forceinline__ ulonglong4 rrr_load_hdr(const RrrImg &v, u64 sb) {
    ulonglong4 h;
    const ulonglong2 *p = reinterpret_cast<const ulonglong2 *>(v.hdr + sb);
    ulonglong2 a, b;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(a.x), "=l"(a.y) : "l"(p));
    asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(b.x), "=l"(b.y) : "l"(p + 1));
    h.x = a.x; h.y = a.y; h.z = b.x; h.w = b.y;
    return h;
}

// ones in [0, i] and bit i; i < n_bits
__device__ __forceinline__ u64 rrr_rank_incl_bit(const RrrImg &v, const RrrTables &T, u64 i, bool *bit) {
    u64 blk = i / kRrrBlock;
    u32 o = u32(i - blk * kRrrBlock);
    u64 sb = blk / kRrrPerSb;
    u32 j = u32(blk - sb * kRrrPerSb);
    ulonglong4 h = rrr_load_hdr(v, sb);
    u64 ones = h.x, bp = h.y;
#pragma unroll
    for (u32 t = 0; t < kRrrPerSb; ++t) {
        if (t < j) {
            u32 c = hdr_class(h, t);
            ones += c;
            bp += T.space[c];
        }
    }
    u32 k = hdr_class(h, j);
    u64 N = read_bits(v.offs, bp, T.space[k]);
    u32 above = rrr_ones_above(T, k, N, o, bit);
    return ones + k - above;
}

// position of the q-th (0-based) one; q < n_ones
__device__ __forceinline__ u64 rrr_select(const RrrImg &v, const RrrTables &T, u64 q) {
    u64 si = q >> v.s1;
    u64 lo = __ldg(v.sel1 + si), hi = __ldg(v.sel1 + si + 1);
    while (lo < hi) {
        u64 mid = (lo + hi + 1) >> 1;
        u64 before = __ldg(reinterpret_cast<const u64 *>(v.hdr + mid));
        if (before <= q) lo = mid; else hi = mid - 1;
    }
    ulonglong4 h = rrr_load_hdr(v, lo);
    u32 r = u32(q - h.x);
    u64 bp = h.y;
    u32 j = 0, k = 0;
    for (; j < kRrrPerSb; ++j) {
        k = hdr_class(h, j);
        if (r < k) break;
        r -= k;
        bp += T.space[k];
    }
    u64 N = read_bits(v.offs, bp, T.space[k]);
    return (lo * kRrrPerSb + j) * kRrrBlock + rrr_select_in_block(T, k, N, r);
}

// ---------------------------------------------------------------------------------------------
// construction from plain device words
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ u64 plain_block63(const u64 *plain, u64 n_words, u64 n_bits, u64 blk) {
    u64 bp = blk * kRrrBlock;
    if (bp >= n_bits) return 0;
    u64 w = bp >> 6;
    u32 o = u32(bp & 63);
    u64 x = plain[w] >> o;
    if (o > 1 && w + 1 < n_words) x |= plain[w + 1] << (64 - o);
    return x & ((1ull << kRrrBlock) - 1ull);   // padding bits of the plain words are zero
}

__global__ void rrr_count_kernel(const u64 *plain, u64 n_words, u64 n_bits, u64 n_sb, u64 *sb_ones, u64 *sb_bits) {
    u64 sb = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (sb >= n_sb) return;
    u32 ones = 0, bits = 0;
    for (u32 j = 0; j < kRrrPerSb; ++j) {
        u32 k = __popcll(plain_block63(plain, n_words, n_bits, sb * kRrrPerSb + j));
        ones += k;
        bits += g_rrr_tables.space[k];
    }
    sb_ones[sb] = ones;
    sb_bits[sb] = bits;
}

__global__ void rrr_encode_kernel(const u64 *plain, u64 n_words, u64 n_bits, u64 n_sb, const u64 *ones_before,
                                  const u64 *bits_before, ulonglong4 *hdr, u64 *offs) {
    u64 sb = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (sb >= n_sb) return;
    ulonglong4 h;
    h.x = ones_before[sb];
    h.y = bits_before[sb];
    h.z = 0; h.w = 0;
    u64 bp = h.y;
    for (u32 j = 0; j < kRrrPerSb; ++j) {
        u64 x = plain_block63(plain, n_words, n_bits, sb * kRrrPerSb + j);
        u32 k = __popcll(x);
        if (j < 8) h.z |= u64(k) << (j * 8); else h.w |= u64(k) << ((j - 8) * 8);
        u32 sp = g_rrr_tables.space[k];
        if (sp) {
            u64 N = 0;
            u32 t = 0;
            while (x) {
                u32 c = __ffsll((long long)x) - 1;
                x &= x - 1;
                ++t;
                N += g_rrr_tables.binom[c][t];
            }
            u32 o = u32(bp & 63);
            atomicOr(offs + (bp >> 6), N << o);
            if (o + sp > 64) atomicOr(offs + (bp >> 6) + 1, N >> (64 - o));
            bp += sp;
        }
    }
    hdr[sb] = h;
}

__global__ void rrr_sample_kernel(const ulonglong4 *hdr, u64 n_sb, u32 s, u32 *sel, u64 n_samples, u64 total) {
    u64 sb = u64(blockIdx.x) * blockDim.x + threadIdx.x;
    if (sb >= n_sb) return;
    ulonglong4 h = hdr[sb];
    u32 c = 0;
    for (u32 j = 0; j < kRrrPerSb; ++j) c += hdr_class(h, j);
    u64 lo = h.x, hi = h.x + c;
    u64 first = (lo + (1ull << s) - 1) >> s;
    for (u64 j = first; (j << s) < hi; ++j) sel[j] = u32(sb);
    if (sb == n_sb - 1)
        for (u64 j = (total + (1ull << s) - 1) >> s; j <= n_samples; ++j) sel[j] = u32(sb);
}

struct RrrBuf {
    DeviceBuf<ulonglong4> hdr;
    DeviceBuf<u64> offs;
    DeviceBuf<u32> sel1;
    RrrImg img{};
    size_t bytes() const { return hdr.bytes() + offs.bytes() + sel1.bytes(); }
};

inline void rrr_build(RrrBuf *out, const u64 *d_plain, u64 n_bits, cudaStream_t st = nullptr) {
    rrr_init_tables();
    u64 n_words = (n_bits + 63) / 64;
    u64 n_sb = n_bits / kRrrSbBits + 1;
    if (n_sb > 0xFFFFFFF0ull) throw LimitError("bit vector too long");
    DeviceBuf<u64> c_ones, c_bits, b_ones, b_bits;
    c_ones.alloc(n_sb + 1); c_bits.alloc(n_sb + 1); b_ones.alloc(n_sb + 1); b_bits.alloc(n_sb + 1);
    u32 g = u32((n_sb + 127) / 128);
    rrr_count_kernel<<<g, 128, 0, st>>>(d_plain, n_words, n_bits, n_sb, c_ones.p, c_bits.p);
    CUDA_CHECK(cudaMemsetAsync(c_ones.p + n_sb, 0, 8, st));
    CUDA_CHECK(cudaMemsetAsync(c_bits.p + n_sb, 0, 8, st));
    size_t need = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, need, c_ones.p, b_ones.p, n_sb + 1, st);
    DeviceBuf<unsigned char> tmp;
    tmp.alloc(need + 16);
    cub::DeviceScan::ExclusiveSum(tmp.p, need, c_ones.p, b_ones.p, n_sb + 1, st);
    cub::DeviceScan::ExclusiveSum(tmp.p, need, c_bits.p, b_bits.p, n_sb + 1, st);
    u64 tot[2] = {0, 0};
    CUDA_CHECK(cudaMemcpyAsync(&tot[0], b_ones.p + n_sb, 8, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(&tot[1], b_bits.p + n_sb, 8, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    u64 off_words = (tot[1] + 63) / 64 + 2;
    out->offs.alloc(off_words);
    CUDA_CHECK(cudaMemsetAsync(out->offs.p, 0, off_words * 8, st));
    out->hdr.alloc(n_sb);
    rrr_encode_kernel<<<g, 128, 0, st>>>(d_plain, n_words, n_bits, n_sb, b_ones.p, b_bits.p, out->hdr.p, out->offs.p);
    CUDA_CHECK(cudaGetLastError());
    RrrImg &im = out->img;
    im.hdr = out->hdr.p; im.offs = out->offs.p;
    im.n_bits = n_bits; im.n_ones = tot[0]; im.n_sb = n_sb;
    double per = double(tot[0]) * kRrrSpanSb / double(n_sb);
    u32 s = 0;
    while (s < 16 && double(2ull << s) <= per) ++s;
    im.s1 = s;
    u64 ns = (tot[0] >> s) + 1;
    out->sel1.alloc(ns + 1);
    CUDA_CHECK(cudaMemsetAsync(out->sel1.p, 0, (ns + 1) * 4, st));
    rrr_sample_kernel<<<g, 128, 0, st>>>(out->hdr.p, n_sb, s, out->sel1.p, ns, tot[0]);
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaStreamSynchronize(st));
    im.sel1 = out->sel1.p;
}

}  // namespace ggq
