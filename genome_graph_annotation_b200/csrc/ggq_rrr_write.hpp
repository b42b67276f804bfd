// Host-side WRITER of the reference's stream formats — the inverse of ggq_format.hpp — so that a matrix that lives
// only as a device image (built on the device, or loaded without keeping its stream) can be serialized as the
// reference would have serialized it: BRWT::serialize (annotation/hierarchical_annotation/BRWT.cpp:113-128) =
// RangePartition (common/utils.cpp:717-725), rrr_vector<63> (sdsl rrr_vector.hpp:350-362), child count, children.
//
// The rrr_vector<63, int_vector<>, 32> arrays are produced by the construction rule of sdsl
// (rrr_vector.hpp:145-238): per 63-bit block its class (popcount, 6 bits) and its offset in the reference's
// combinatorial numbering (rrr_helper.hpp:295-313: walking the block from bit 0, a set bit at distance i from the
// start with k ones still to place adds C(62 - i, k)); per superblock of 32 blocks a pointer into the offsets, a rank
// sample and an inversion flag (a full superblock with more than 16 blocks of class > 31 stores 63 - class).
// Byte-identical to the reference's own serialization of the same bits (tests/test_oracle_cpu.py).
#pragma once
#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include "ggq_format.hpp"

namespace ggq {

// Where the bytes go: a growing string, or a caller's write callback (returns the bytes it took; fewer = failure).
struct ByteSink {
    std::string *out = nullptr;
    size_t (*write)(void *ctx, const void *src, size_t n) = nullptr;
    void *ctx = nullptr;
    uint64_t total = 0;
    void raw(const void *p, size_t n) {
        if (!n) return;
        total += n;
        if (out) out->append(static_cast<const char *>(p), n);
        else if (!write || write(ctx, p, n) != n) throw std::runtime_error("serialize: the output stream refused the data");
    }
    void u64le(uint64_t v) { raw(&v, 8); }
    void u64be(uint64_t v) {
        unsigned char b[8];
        for (int i = 0; i < 8; ++i) b[i] = (unsigned char)(v >> (56 - 8 * i));
        raw(b, 8);
    }
    void u8(uint8_t v) { raw(&v, 1); }
};

// sdsl::int_vector<0> under construction: `count` values of `width` bits, LSB-first in u64 words
struct PackedInts {
    uint64_t count = 0;
    uint8_t width = 1;
    std::vector<uint64_t> words;
    PackedInts(uint64_t n, uint8_t w) : count(n), width(w), words((n * w + 63) / 64, 0) {}
    void set(uint64_t i, uint64_t v) { put_bits(i * width, v, width); }
    uint64_t get(uint64_t i) const {
        const uint64_t bp = i * width, w = bp >> 6;
        const unsigned o = bp & 63;
        uint64_t x = words[w] >> o;
        if (o + width > 64) x |= words[w + 1] << (64 - o);
        return width == 64 ? x : x & ((uint64_t(1) << width) - 1);
    }
    // (the target bits are zero: values are only ever set once, or overwritten through clear_bits first)
    void put_bits(uint64_t bp, uint64_t v, unsigned len) {
        if (!len) return;
        const uint64_t w = bp >> 6;
        const unsigned o = bp & 63;
        words[w] |= v << o;
        if (o + len > 64) words[w + 1] |= v >> (64 - o);
    }
    void clear(uint64_t i) {
        const uint64_t bp = i * width, w = bp >> 6;
        const unsigned o = bp & 63;
        const uint64_t mask = width == 64 ? ~uint64_t(0) : (uint64_t(1) << width) - 1;
        words[w] &= ~(mask << o);
        if (o + width > 64) words[w + 1] &= ~(mask >> (64 - o));
    }
    void write(ByteSink &s, bool with_width) const {   // int_vector<0> / bit_vector (no width byte)
        s.u64le(count * width);
        if (with_width) s.u8(width);
        s.raw(words.data(), words.size() * 8);
    }
};

inline unsigned hi_bit(uint64_t x) { return x ? 63u - unsigned(__builtin_clzll(x)) : 0u; }   // sdsl::bits::hi

// `len` (<= 63) bits of the plain vector starting at bit `pos` (words are zero-padded)
inline uint64_t plain_bits(const uint64_t *w, uint64_t n_words, uint64_t pos, unsigned len) {
    const uint64_t wi = pos >> 6;
    const unsigned o = pos & 63;
    uint64_t x = wi < n_words ? w[wi] >> o : 0;
    if (o + len > 64 && wi + 1 < n_words) x |= w[wi + 1] << (64 - o);
    return x & ((uint64_t(1) << len) - 1);
}

// rrr_helper<63>::bin_to_nr (rrr_helper.hpp:295-313): the block's offset among the blocks of its class
// (one step per SET bit — position by count-trailing-zeros — instead of one unpredictable branch per position: the
// additions are independent of each other, ~20 ns per dense block instead of ~300)
inline uint64_t rrr63_block_number(uint64_t bin, const Binomial63 &B) {
    uint64_t nr = 0;
    unsigned k = unsigned(__builtin_popcountll(bin));
    while (bin) {
        const unsigned i = unsigned(__builtin_ctzll(bin));   // a set bit at distance i from the start, k ones still to place
        nr += B.c[62 - i][k];                                // (C(n, k > n) = 0: a full block numbers 0)
        --k;
        bin &= bin - 1;
    }
    return nr;
}

// Runs fn(t, n_threads) on n_threads host threads (1: inline).
template <typename F>
inline void rrr_parallel(unsigned n_threads, F fn) {
    if (n_threads <= 1) { fn(0u, 1u); return; }
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < n_threads; ++t) pool.emplace_back([&, t] { fn(t, n_threads); });
    for (std::thread &th : pool) th.join();
}

// rrr_vector<63>(bv).serialize(out): words = the plain bits, LSB first, (n_bits + 63) / 64 of them, zero-padded.
// Three steps so that long vectors are encoded by several host threads: (1) every block's class, in parallel;
// (2) one cheap sequential sweep over the superblocks in sdsl's order of events — inversion flag, offset pointer, rank
// sample; (3) stored classes and offsets, in parallel (a superblock's 32 six-bit classes fill exactly three words; the
// other arrays are written with atomic ORs into zeroed words).
inline void write_rrr63(ByteSink &s, const uint64_t *words, uint64_t n_bits) {
    const unsigned BS = 63, K = 32;
    const Binomial63 &B = Binomial63::get();
    const uint64_t n_words = (n_bits + 63) / 64;
    const uint64_t n_blocks = (n_bits + BS) / BS;   // a dummy block at the end when n_bits % 63 == 0
    const uint64_t n_real = (n_bits + BS - 1) / BS; // blocks that hold bits (the last one possibly partial)
    const uint64_t n_sb = (n_blocks + K - 1) / K;
    unsigned n_thr = 1;
    if (n_blocks >= (uint64_t(1) << 16)) {
        n_thr = std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
        if (const char *e = std::getenv("GGQ_LOAD_THREADS")) n_thr = unsigned(std::max(1, std::atoi(e)));
    }
    auto block_bits = [&](uint64_t b) {
        const uint64_t pos = b * BS;
        return plain_bits(words, n_words, pos, unsigned(std::min<uint64_t>(BS, n_bits - pos)));
    };
    // (1) classes
    std::vector<uint8_t> cls(n_blocks, 0);
    rrr_parallel(n_thr, [&](unsigned t, unsigned T) {
        for (uint64_t b = n_real * t / T; b < n_real * (t + 1) / T; ++b) cls[b] = uint8_t(__builtin_popcountll(block_bits(b)));
    });
    // (2) per superblock: inverted?, first offset bit, ones before.  sdsl opens a superblock when it REACHES its first
    // block, so a superblock that starts at the dummy block is never opened (its pointer and sample stay 0), and one
    // that starts at the partial last block is opened without the inversion test.
    std::vector<uint8_t> inv(n_sb, 0), opened(n_sb, 0);
    std::vector<uint64_t> at(n_sb + 1, 0), before(n_sb + 1, 0);
    uint64_t bit_at = 0, ones = 0;
    for (uint64_t sb = 0; sb < n_sb; ++sb) {
        const uint64_t b0 = sb * K, b1 = std::min<uint64_t>(b0 + K, n_blocks);
        at[sb] = bit_at;
        before[sb] = ones;
        opened[sb] = b0 < n_real;
        const bool starts_full = (b0 + 1) * BS <= n_bits;   // its first block is a full one
        if (opened[sb] && starts_full && b0 + K <= n_blocks) {
            unsigned dense = 0;
            for (uint64_t b = b0; b < b1; ++b) dense += cls[b] > BS / 2;
            inv[sb] = dense > K / 2;
        }
        for (uint64_t b = b0; b < b1 && b < n_real; ++b) { bit_at += B.space[cls[b]]; ones += cls[b]; }
    }
    at[n_sb] = bit_at;
    before[n_sb] = ones;
    PackedInts bt(n_blocks, 6);
    PackedInts btnr(std::max<uint64_t>(bit_at, 64), 1);
    PackedInts btnrp(n_sb, uint8_t(hi_bit(bit_at) + 1));
    PackedInts rank(n_sb + ((n_bits % (uint64_t(K) * BS)) > 0 ? 1 : 0), uint8_t(hi_bit(ones) + 1));
    PackedInts invert(n_sb, 1);
    auto or_bits = [](PackedInts &p, uint64_t bp, uint64_t v, unsigned len) {   // concurrent writers share words
        if (!len || !v) return;
        const uint64_t w = bp >> 6;
        const unsigned o = bp & 63;
        __atomic_fetch_or(&p.words[w], v << o, __ATOMIC_RELAXED);
        if (o + len > 64) __atomic_fetch_or(&p.words[w + 1], v >> (64 - o), __ATOMIC_RELAXED);
    };
    // (3) stored classes, offsets, pointers, samples, flags
    rrr_parallel(n_thr, [&](unsigned t, unsigned T) {
        for (uint64_t sb = n_sb * t / T; sb < n_sb * (t + 1) / T; ++sb) {
            const uint64_t b0 = sb * K, b1 = std::min<uint64_t>(b0 + K, n_blocks);
            if (opened[sb]) {
                or_bits(btnrp, sb * btnrp.width, at[sb], btnrp.width);
                or_bits(rank, sb * rank.width, before[sb], rank.width);
                if (inv[sb]) or_bits(invert, sb, 1, 1);
            }
            uint64_t p = at[sb];
            for (uint64_t b = b0; b < b1; ++b) {
                const unsigned stored = inv[sb] ? BS - cls[b] : cls[b];
                or_bits(bt, b * 6, stored, 6);
                if (b >= n_real) continue;
                const unsigned sp = B.space[cls[b]];   // space[k] == space[63 - k]
                if (sp) or_bits(btnr, p, rrr63_block_number(block_bits(b), B), sp);
                p += sp;
            }
        }
    });
    rank.clear(rank.count - 1);
    rank.set(rank.count - 1, ones);   // "for technical reasons": the total closes the samples
    s.u64le(n_bits);
    bt.write(s, true);
    btnr.write(s, false);
    btnrp.write(s, true);
    rank.write(s, true);
    invert.write(s, false);
}

// serialize_number_vector<uint32_t> (common/serialization.cpp:38-50): an sdsl::int_vector<> of width 32
inline void write_u32_vector(ByteSink &s, const std::vector<uint32_t> &v) {
    PackedInts p(v.size(), 32);
    for (size_t i = 0; i < v.size(); ++i) p.set(i, v[i]);
    p.write(s, true);
}

}  // namespace ggq
