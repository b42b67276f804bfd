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
#include <string>
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
inline uint64_t rrr63_block_number(uint64_t bin, const Binomial63 &B) {
    if (bin == 0 || bin == ((uint64_t(1) << 63) - 1)) return 0;
    uint64_t nr = 0;
    unsigned k = unsigned(__builtin_popcountll(bin)), nn = 63;
    while (bin) {
        if (bin & 1) { nr += B.c[nn - 1][k]; --k; }
        bin >>= 1;
        --nn;
    }
    return nr;
}

// rrr_vector<63>(bv).serialize(out): words = the plain bits, LSB first, (n_bits + 63) / 64 of them, zero-padded
inline void write_rrr63(ByteSink &s, const uint64_t *words, uint64_t n_bits) {
    const unsigned BS = 63, K = 32;
    const Binomial63 &B = Binomial63::get();
    const uint64_t n_words = (n_bits + 63) / 64;
    const uint64_t n_blocks = (n_bits + BS) / BS;   // a dummy block at the end when n_bits % 63 == 0
    const uint64_t n_sb = (n_blocks + K - 1) / K;
    PackedInts bt(n_blocks, 6);
    // (1) block classes, total offset bits, total ones
    uint64_t pos = 0, i = 0, btnr_bits = 0, ones = 0;
    while (pos + BS <= n_bits) {
        const unsigned x = unsigned(__builtin_popcountll(plain_bits(words, n_words, pos, BS)));
        bt.set(i++, x);
        ones += x;
        btnr_bits += B.space[x];
        pos += BS;
    }
    if (pos < n_bits) {
        const unsigned x = unsigned(__builtin_popcountll(plain_bits(words, n_words, pos, unsigned(n_bits - pos))));
        bt.set(i++, x);
        ones += x;
        btnr_bits += B.space[x];
    }
    PackedInts btnr(std::max<uint64_t>(btnr_bits, 64), 1);
    PackedInts btnrp(n_sb, uint8_t(hi_bit(btnr_bits) + 1));
    PackedInts rank(n_sb + ((n_bits % (uint64_t(K) * BS)) > 0 ? 1 : 0), uint8_t(hi_bit(ones) + 1));
    PackedInts invert(n_sb, 1);
    // (2) offsets, pointers, rank samples, inversion flags — in sdsl's order of events
    pos = 0; i = 0;
    uint64_t at = 0, sum = 0;
    bool inv = false;
    auto open_superblock = [&](bool may_invert) {
        btnrp.set(i / K, at);
        rank.set(i / K, sum);
        inv = false;
        if (may_invert && i + K <= n_blocks) {
            unsigned dense = 0;
            for (uint64_t j = i; j < i + K; ++j) dense += bt.get(j) > BS / 2;
            if (dense > K / 2) {
                invert.set(i / K, 1);
                for (uint64_t j = i; j < i + K; ++j) { const uint64_t v = bt.get(j); bt.clear(j); bt.set(j, BS - v); }
                inv = true;
            }
        }
    };
    while (pos + BS <= n_bits) {
        if (i % K == 0) open_superblock(true);
        const unsigned x = unsigned(bt.get(i++));
        const unsigned sp = B.space[x];
        sum += inv ? BS - x : x;
        if (sp) btnr.put_bits(at, rrr63_block_number(plain_bits(words, n_words, pos, BS), B), sp);
        at += sp;
        pos += BS;
    }
    if (pos < n_bits) {   // the last, partial block (never starts an inverted superblock)
        if (i % K == 0) open_superblock(false);
        const unsigned x = unsigned(bt.get(i++));
        const unsigned sp = B.space[x];
        sum += inv ? BS - x : x;
        if (sp) btnr.put_bits(at, rrr63_block_number(plain_bits(words, n_words, pos, unsigned(n_bits - pos)), B), sp);
        at += sp;
    }
    rank.clear(rank.count - 1);
    rank.set(rank.count - 1, sum);   // "for technical reasons": the total closes the samples
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
