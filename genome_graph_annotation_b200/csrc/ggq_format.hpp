// Host-side readers for the reference's on-disk formats (no sdsl / libmaus2 dependency).
//
// What is parsed, and where the reference writes it:
//   * u64 big-endian numbers            libmaus2 NumberSerialisation (common/serialization.cpp:27-36)
//   * sdsl::int_vector<> / bit_vector   sdsl/int_vector.hpp:593-609,1565-1598 (LE u64 bit size, [u8 width], LE words)
//   * utils::RangePartition             common/utils.cpp:702-725
//   * sdsl::rrr_vector<63>              sdsl/rrr_vector.hpp:350-373 (m_size, bt, btnr, btnrp, rank, invert)
//   * sdsl::sd_vector<> + inverted byte sdsl/sd_vector.hpp:404-432, common/bit_vector.cpp:725-756
//   * bit_vector_stat                   common/bit_vector.cpp:346-385
//   * BRWT node, pre-order              annotation/hierarchical_annotation/BRWT.cpp:87-128
//   * ColMajorCompressed                annotation/column_compressed/column_major.cpp:58-88
//
// Everything is decoded to PLAIN bits here; the device layout is built from those
// (ggq_engine.cu).  Errors are reported by throwing ParseError; the C ABI maps it to
// GGQ_ERR_FORMAT, and the BinaryMatrix adapter maps that to load() == false.
#pragma once

#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

// ggq_rrr_simd.cpp (host compiler only): the AVX-512 form of the block walk.  Weak: a program that includes this header
// without linking that file simply uses the scalar walk.
extern "C" int ggq_rrr_simd_available() __attribute__((weak));
extern "C" int ggq_rrr_walk16(const uint64_t (*walk)[64], const unsigned *k, const uint64_t *nr, uint64_t *bits) __attribute__((weak));

namespace ggq {

struct ParseError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// Pull-style byte source: memory buffer or a caller-provided callback (std::istream in the adapter).
class Reader {
  public:
    typedef size_t (*read_fn)(void *ctx, void *dst, size_t n);

    Reader(const uint8_t *buf, size_t len) : buf_(buf), len_(len) {}
    Reader(read_fn fn, void *ctx) : fn_(fn), ctx_(ctx) {}

    void read(void *dst, size_t n) {
        if (!n) return;
        if (fn_) {
            if (fn_(ctx_, dst, n) != n) throw ParseError("unexpected end of stream");
        } else {
            if (n > len_ - pos_) throw ParseError("unexpected end of buffer");
            std::memcpy(dst, buf_ + pos_, n);
        }
        pos_ += n;
    }
    // bytes still available if known (memory source), otherwise "unknown" = huge
    uint64_t remaining() const { return fn_ ? ~uint64_t(0) : len_ - pos_; }
    uint64_t consumed() const { return pos_; }

    uint8_t u8() { uint8_t v; read(&v, 1); return v; }
    uint64_t u64le() { uint64_t v; read(&v, 8); return v; }  // host is little-endian (x86-64)
    uint64_t u64be() {
        uint8_t b[8];
        read(b, 8);
        uint64_t v = 0;
        for (int i = 0; i < 8; ++i) v = (v << 8) | b[i];
        return v;
    }

  private:
    const uint8_t *buf_ = nullptr;
    size_t len_ = 0, pos_ = 0;
    read_fn fn_ = nullptr;
    void *ctx_ = nullptr;
};

// sdsl::int_vector<0> (has_width = true) or int_vector<t_width> with fixed width (bit_vector: 1, <64>: 64).
struct IntVector {
    uint64_t bit_size = 0;
    uint8_t width = 1;
    std::vector<uint64_t> words;

    uint64_t size() const { return width ? bit_size / width : 0; }

    // value of element i (width <= 64); same bit addressing as sdsl::bits::read_int
    uint64_t get(uint64_t i) const {
        if (i >= size()) throw ParseError("int_vector index out of range");   // (also keeps i * width from wrapping)
        return get_bits(i * width, width);
    }

    uint64_t get_bits(uint64_t bit_pos, unsigned len) const {
        if (!len) return 0;
        uint64_t w = bit_pos >> 6;
        unsigned off = bit_pos & 63;
        uint64_t v = words[w] >> off;
        if (off + len > 64) v |= words[w + 1] << (64 - off);
        return len == 64 ? v : v & ((uint64_t(1) << len) - 1);
    }
    bool bit(uint64_t i) const { return (words[i >> 6] >> (i & 63)) & 1; }

    void load(Reader &in, bool has_width, uint8_t fixed_width = 1) {
        bit_size = in.u64le();
        width = has_width ? in.u8() : fixed_width;
        uint64_t n_words = (bit_size + 63) >> 6;
        if (n_words < (bit_size >> 6)) throw ParseError("int_vector size overflow");
        if (n_words > in.remaining() / 8 && in.remaining() != ~uint64_t(0))
            throw ParseError("int_vector larger than the stream");
        if (n_words > (uint64_t(1) << 40)) throw ParseError("int_vector implausibly large");
        words.assign(n_words + 1, 0);  // +1 pad word so that get_bits may peek one word ahead
        in.read(words.data(), n_words * 8);
        if (has_width && bit_size && (width == 0 || width > 64)) throw ParseError("bad int_vector width");
    }
};

// Plain bits of one vector, as u32 words (bit i -> word i/32, bit i%32), zero padded.
struct PlainBits {
    uint64_t size = 0;  // in bits
    uint64_t ones = 0;
    std::vector<uint32_t> w;

    void resize(uint64_t n_bits) {
        size = n_bits;
        w.assign((n_bits + 31) / 32 + 4, 0);
    }
    void set(uint64_t i) { w[i >> 5] |= uint32_t(1) << (i & 31); }
    bool get(uint64_t i) const { return (w[i >> 5] >> (i & 31)) & 1; }
    // OR `len` (<= 64) bits of v at bit position pos
    void or_bits(uint64_t pos, uint64_t v, unsigned len) {
        if (!len) return;
        if (len < 64) v &= (uint64_t(1) << len) - 1;
        uint64_t wi = pos >> 5;
        unsigned off = pos & 31;
        w[wi] |= uint32_t(v << off);
        v >>= (32 - off);
        w[wi + 1] |= uint32_t(v);
        if (off + len > 64) w[wi + 2] |= uint32_t(v >> 32);
    }
    void count() {
        ones = 0;
        for (uint32_t x : w) ones += __builtin_popcount(x);
    }
};

// ---------------------------------------------------------------------------------------
// rrr_vector<63, int_vector<>, 32>
// ---------------------------------------------------------------------------------------
struct Binomial63 {
    uint64_t c[64][64];   // c[n][k], n,k <= 63
    uint64_t walk[63][64];  // c[n][k] with walk[n][0] = 2^63 ("never taken"): the table of decode_batch
    uint8_t space[64];    // bits used to store the block number of a block with k ones (sdsl rrr_helper.hpp:267-272)
    Binomial63() {
        for (int n = 0; n < 64; ++n)
            for (int k = 0; k < 64; ++k)
                c[n][k] = (k == 0) ? 1 : 0;
        for (int n = 1; n < 64; ++n)
            for (int k = 1; k <= n; ++k)
                c[n][k] = c[n - 1][k - 1] + c[n - 1][k];
        for (int k = 0; k < 64; ++k) {
            uint64_t b = c[63][k];
            space[k] = (b == 1) ? 0 : uint8_t(64 - __builtin_clzll(b));
        }
        for (int n = 0; n < 63; ++n)
            for (int k = 0; k < 64; ++k) walk[n][k] = k == 0 ? (uint64_t(1) << 63) : c[n][k];
    }
    static const Binomial63 &get() {
        static const Binomial63 t;
        return t;
    }
    // The same walk for N blocks at once, without data-dependent branches.  One block's walk is a chain of dependent
    // steps (k -> table entry -> compare -> k), and on a BRWT's vectors — dense below the root: a child is as long as
    // its parent has ones — the branch "bit set?" is a coin flip: the scalar loop spends 300-550 ns per block.  N
    // independent chains in one loop fill the core's issue slots instead (~100 ns per block).  Walking on to k == 0
    // (C(n, 1) = n reproduces the closed form of the last one, C(n, k > n) = 0 forces the remaining ones; walk[n][0] is
    // a value no offset reaches, so a finished lane takes nothing more) leaves nr == 0 for every valid (k, nr);
    // anything else is a corrupt block.
    template <int N>
    void decode_batch(const unsigned *k_in, const uint64_t *nr_in, unsigned n, uint64_t *bits_out) const {
        uint64_t k[N], nr[N], bits[N];
        for (int j = 0; j < N; ++j) {
            const bool live = unsigned(j) < n;
            k[j] = live ? k_in[j] : 0;
            nr[j] = live ? nr_in[j] : 0;
            bits[j] = 0;
        }
        for (unsigned i = 0; i < 63; ++i) {
            const uint64_t *row = walk[62 - i];
            const uint64_t bit = uint64_t(1) << i;
            auto lane = [&](int j) {
                const uint64_t cv = row[k[j]];
                const uint64_t m = 0 - uint64_t(nr[j] >= cv);   // all ones: the bit is set
                nr[j] -= cv & m;
                k[j] += m;
                bits[j] |= bit & m;
            };
            static_assert(N == 8, "the lanes are written out: keep them in registers");
            lane(0); lane(1); lane(2); lane(3); lane(4); lane(5); lane(6); lane(7);
        }
        for (unsigned j = 0; j < n; ++j) {
            if (k[j] != 0 || nr[j] != 0) throw ParseError("corrupt rrr block");
            bits_out[j] = bits[j];
        }
    }

    // Inverse of sdsl's bin_to_nr (rrr_helper.hpp:295-312): bit i of the block is set iff
    // nr >= C(62 - i, k) at the moment i is visited.
    uint64_t decode(unsigned k, uint64_t nr) const {
        if (k == 0) return 0;
        if (k == 63) return (uint64_t(1) << 63) - 1;
        uint64_t bits = 0;
        unsigned nn = 63;
        unsigned i = 0;
        while (k > 1) {
            if (nn == 0) throw ParseError("corrupt rrr block");
            if (nr >= c[nn - 1][k]) {
                nr -= c[nn - 1][k];
                --k;
                bits |= uint64_t(1) << i;
            }
            --nn;
            ++i;
        }
        if (nr > 62) throw ParseError("corrupt rrr block");
        uint64_t last = 62 - nr;  // position of the final one: n - nr - 1
        bits |= uint64_t(1) << last;
        return bits;
    }
};

// The serialized arrays of one rrr_vector<63>, undecoded: about the size they have in the stream.  A BRWT keeps
// these per node until the node is uploaded (lazy decode: the host never holds the plain bits of more than one node).
struct RrrRaw {
    uint64_t m_size = 0;
    IntVector bt, btnr, btnrp, invert;
    bool empty = true;   // default-constructed rrr_vector (empty matrix) / nothing read
    void clear() { *this = RrrRaw(); }
};

inline void read_rrr63_raw(Reader &in, RrrRaw *r) {
    const unsigned BS = 63, K = 32;
    r->m_size = in.u64le();
    IntVector rank;
    r->bt.load(in, true);
    r->btnr.load(in, false, 1);
    r->btnrp.load(in, true);
    rank.load(in, true);
    r->invert.load(in, false, 1);
    r->empty = r->m_size == 0 && r->bt.bit_size == 0;
    if (r->empty) return;
    const uint64_t n_blocks = (r->m_size + BS) / BS;  // includes the trailing dummy/partial block
    if (r->bt.width < 6 || r->bt.size() != n_blocks) throw ParseError("rrr: bad block-type array");
    const uint64_t n_sb = (n_blocks + K - 1) / K;
    if (r->invert.bit_size < n_sb || r->btnrp.size() < n_sb) throw ParseError("rrr: bad superblock arrays");
}

// max_threads: 0 = as many as the vector's length warrants (below); 1 = the calling thread only (the caller runs
// several decodes side by side: decode_nodes_parallel, ggq_tree.hpp)
inline void decode_rrr63(const RrrRaw &raw, PlainBits *out, unsigned max_threads = 0) {
    const unsigned BS = 63, K = 32;
    if (raw.empty) { out->resize(0); return; }
    const uint64_t m_size = raw.m_size;
    const IntVector &bt = raw.bt, &btnr = raw.btnr, &btnrp = raw.btnrp, &invert = raw.invert;
    const uint64_t n_blocks = (m_size + BS) / BS;
    const uint64_t n_sb = (n_blocks + K - 1) / K;
    const Binomial63 &B = Binomial63::get();
    static const bool simd = [] {   // GGQ_RRR_SIMD=0: the scalar walk (tests run both)
        const char *e = std::getenv("GGQ_RRR_SIMD");
        return !(e && *e == '0') && ggq_rrr_simd_available && ggq_rrr_walk16 && ggq_rrr_simd_available();
    }();
    out->resize(m_size);
    // Superblocks (32 blocks = 2016 bits = 63 u32 words) are independent: each starts at its own offset pointer and
    // on a word boundary of the output, so large vectors are decoded by several host threads.
    auto decode_range = [&](uint64_t sb0, uint64_t sb1) {
        constexpr int kBatch = 16;   // blocks walked together: one AVX-512 walk, or two scalar walks of eight
        unsigned qk[kBatch], qlen[kBatch], nq = 0;
        uint64_t qnr[kBatch], qstart[kBatch], qbits[kBatch];
        auto flush = [&] {
            if (!nq) return;
            if (simd) {
                for (unsigned j = nq; j < kBatch; ++j) { qk[j] = 0; qnr[j] = 0; }   // idle lanes: k = 0, nr = 0 is a valid end state
                if (!ggq_rrr_walk16(B.walk, qk, qnr, qbits)) throw ParseError("corrupt rrr block");
            } else {
                B.decode_batch<8>(qk, qnr, std::min(nq, 8u), qbits);
                if (nq > 8) B.decode_batch<8>(qk + 8, qnr + 8, nq - 8, qbits + 8);
            }
            for (unsigned j = 0; j < nq; ++j) {
                if (qlen[j] < BS && (qbits[j] >> qlen[j])) throw ParseError("rrr: bits beyond the end");
                out->or_bits(qstart[j], qbits[j], qlen[j]);
            }
            nq = 0;
        };
        // (bt holds n_blocks entries of bt.width >= 6 bits and invert / btnrp n_sb entries: checked by read_rrr63_raw)
        const unsigned btw = bt.width;
        uint64_t pos = 0;
        for (uint64_t sb = sb0; sb < sb1; ++sb) {
            const uint64_t b0 = sb * K, b1 = std::min<uint64_t>(n_blocks, b0 + K);
            // The dummy block sdsl appends when m_size is a multiple of 63 holds nothing — and when it opens a superblock
            // (m_size a multiple of 2016) that superblock was never "reached" by the construction loop: its pointer and
            // rank sample are 0 (rrr_vector.hpp:196-238), which is not a pointer running backwards.
            if (b0 * BS >= m_size) break;
            const uint64_t p = btnrp.get(sb);   // file-controlled: must not run backwards or past the offsets
            if (p > btnr.bit_size || (sb > sb0 && p < pos) || (sb && p < btnrp.get(sb - 1)))
                throw ParseError("rrr: superblock pointer out of range");
            pos = p;
            const bool inverted = invert.bit(sb);
            for (uint64_t b = b0; b < b1; ++b) {
                const uint64_t start = b * BS;
                if (start >= m_size) break;   // the dummy block
                const unsigned stored = unsigned(bt.get_bits(b * btw, btw));
                if (stored > BS) throw ParseError("rrr: block type out of range");
                const unsigned sp = B.space[stored];  // space[k] == space[63-k]
                const unsigned k = inverted ? BS - stored : stored;
                const unsigned len = unsigned(std::min<uint64_t>(BS, m_size - start));
                if (sp) {
                    if (pos > btnr.bit_size || sp > btnr.bit_size - pos) throw ParseError("rrr: btnr overrun");
                    const uint64_t nr = btnr.get_bits(pos, sp);
                    if (k <= 2) {   // (sparse vectors: one or two ones — the scalar walk ends early)
                        const uint64_t bits = B.decode(k, nr);
                        if (len < BS && (bits >> len)) throw ParseError("rrr: bits beyond the end");
                        out->or_bits(start, bits, len);
                    } else {
                        qk[nq] = k; qnr[nq] = nr; qstart[nq] = start; qlen[nq] = len;
                        if (++nq == kBatch) flush();
                    }
                    pos += sp;
                } else if (k == BS) {
                    if (len < BS) throw ParseError("rrr: bits beyond the end");
                    out->or_bits(start, ~uint64_t(0), len);
                }
            }
        }
        flush();
    };
    unsigned n_thr = 1;
    if (n_sb >= (uint64_t(1) << 12)) {
        n_thr = std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
        if (const char *e = std::getenv("GGQ_LOAD_THREADS")) n_thr = std::max(1, std::atoi(e));
    }
    if (max_threads) n_thr = std::min(n_thr, max_threads);
    if (n_thr <= 1) {
        decode_range(0, n_sb);
    } else {
        std::vector<std::thread> pool;
        std::vector<std::string> errors(n_thr);
        for (unsigned t = 0; t < n_thr; ++t)
            pool.emplace_back([&, t] {
                try {
                    decode_range(n_sb * t / n_thr, n_sb * (t + 1) / n_thr);
                } catch (const std::exception &e) {
                    errors[t] = e.what()[0] ? e.what() : "rrr: decode failed";
                }
            });
        for (std::thread &th : pool) th.join();
        for (const std::string &e : errors)
            if (!e.empty()) throw ParseError(e);
    }
    out->count();
}

inline void load_rrr63(Reader &in, PlainBits *out) {
    RrrRaw raw;
    read_rrr63_raw(in, &raw);
    decode_rrr63(raw, out);
}

// ---------------------------------------------------------------------------------------
// sd_vector<> followed by the reference's `inverted` byte  (bit_vector_sd)
// ---------------------------------------------------------------------------------------
inline void skip_select_mcl(Reader &in) {
    uint64_t arg_cnt = in.u64le();
    if (!arg_cnt) return;
    uint64_t sb = (arg_cnt + 4095) >> 12;
    IntVector tmp;
    tmp.load(in, true);  // superblock
    IntVector mini_or_long;
    mini_or_long.load(in, false, 1);
    if (mini_or_long.bit_size && mini_or_long.bit_size < sb) throw ParseError("select_mcl: bad indicator");
    for (uint64_t i = 0; i < sb; ++i)
        tmp.load(in, true);  // long superblock or miniblock — both int_vector<0>
}

struct SdColumn {
    uint64_t size = 0;
    bool inverted = false;            // stored vector is the complement (common/bit_vector.cpp:572-590)
    std::vector<uint64_t> positions;  // ascending positions of the ones of the STORED vector
};

inline void load_sd(Reader &in, SdColumn *out) {
    out->size = in.u64le();
    uint8_t wl = in.u8();
    IntVector low, high;
    low.load(in, true);
    high.load(in, false, 1);
    skip_select_mcl(in);  // select_support_mcl<1>
    skip_select_mcl(in);  // select_support_mcl<0>
    // the trailing byte distinguishes bit_vector_sd from a bare sd_vector (tests/test_bit_vector.cpp:432-461)
    uint8_t inv = in.u8();
    if (inv > 1) throw ParseError("bit_vector_sd: bad inverted flag");
    out->inverted = inv;
    uint64_t m = low.size();
    if (low.bit_size && low.width != wl) throw ParseError("sd: low width mismatch");
    if (wl >= 64) throw ParseError("sd: bad wl");
    out->positions.clear();
    out->positions.reserve(m);
    uint64_t i = 0;
    for (uint64_t hp = 0; hp < high.bit_size && i < m; ++hp) {
        if (high.bit(hp)) {
            uint64_t hv = hp - i;
            uint64_t p = (hv << wl) | (wl ? low.get(i) : 0);
            if (p >= out->size || (i && p <= out->positions.back())) throw ParseError("sd: positions not ascending");
            out->positions.push_back(p);
            ++i;
        }
    }
    if (i != m) throw ParseError("sd: high/low mismatch");
}

// logical bits of a bit_vector_sd
inline void sd_to_plain(const SdColumn &c, PlainBits *out) {
    out->resize(c.size);
    if (c.inverted) {
        for (uint64_t i = 0; i < c.size; ++i) out->set(i);
        for (uint64_t p : c.positions) out->w[p >> 5] &= ~(uint32_t(1) << (p & 31));
    } else {
        for (uint64_t p : c.positions) out->set(p);
    }
    out->count();
}

// bit_vector_stat: plain sdsl::bit_vector + BE popcount + rank_support_v5 + select_support_mcl
// (the supports may be missing: common/bit_vector.cpp:367-380 — tolerated the same way).
inline void load_stat(Reader &in, PlainBits *out) {
    IntVector v;
    v.load(in, false, 1);
    out->resize(v.bit_size);
    for (uint64_t i = 0; i < (v.bit_size + 63) / 64; ++i) {
        out->w[2 * i] = uint32_t(v.words[i]);
        out->w[2 * i + 1] = uint32_t(v.words[i] >> 32);
    }
    if (v.bit_size % 64) {  // clear padding
        uint64_t last = v.bit_size / 64;
        uint64_t mask = (uint64_t(1) << (v.bit_size % 64)) - 1;
        uint64_t x = v.words[last] & mask;
        out->w[2 * last] = uint32_t(x);
        out->w[2 * last + 1] = uint32_t(x >> 32);
    }
    out->count();
    try {
        (void)in.u64be();  // num_set_bits
        IntVector rk;
        rk.load(in, false, 64);
        skip_select_mcl(in);
    } catch (const ParseError &) {
        // supports absent: the reference rebuilds them; nothing to do here
    }
}

// ---------------------------------------------------------------------------------------
// BRWT stream
// ---------------------------------------------------------------------------------------
struct ParsedNode {
    std::vector<std::vector<uint32_t>> partition;  // RangePartition::partition_
    uint64_t n_cols = 0;
    PlainBits bits;                                // nonzero_rows_ (eager parse)
    RrrRaw raw;                                    // ... or its serialized arrays (lazy parse: bits.size is set, w empty)
    std::vector<ParsedNode> children;
};

inline void load_range_partition(Reader &in, ParsedNode *node) {
    uint64_t n_groups = in.u64be();
    if (n_groups > (uint64_t(1) << 32)) throw ParseError("RangePartition: too many groups");
    node->partition.assign(n_groups, {});
    uint64_t total = 0;
    for (auto &g : node->partition) {
        IntVector v;
        v.load(in, true);
        uint64_t k = v.size();
        if (!k) throw ParseError("RangePartition: empty group");  // utils.cpp:655-659
        g.resize(k);
        for (uint64_t i = 0; i < k; ++i) g[i] = uint32_t(v.get(i));
        total += k;
    }
    // initialize_groups_and_ranks (utils.cpp:651-680): must be a permutation of [0, total)
    std::vector<uint8_t> seen(total, 0);
    for (auto &g : node->partition)
        for (uint32_t c : g) {
            if (c >= total || seen[c]) throw ParseError("RangePartition: not a partition");
            seen[c] = 1;
        }
    node->n_cols = total;
}

// lazy: keep every node's vector undecoded (decode_rrr63 at upload time)
inline void load_brwt_node(Reader &in, ParsedNode *node, bool lazy = false, int depth = 0) {
    if (depth > 4096) throw ParseError("BRWT: implausible depth");
    load_range_partition(in, node);
    if (lazy) {
        read_rrr63_raw(in, &node->raw);
        if (node->raw.empty) node->bits.resize(0);   // nothing to decode: an ordinary (empty) vector
        else node->bits.size = node->raw.m_size;
    } else {
        load_rrr63(in, &node->bits);
    }
    uint64_t n_children = in.u64be();
    if (n_children && n_children != node->partition.size())
        throw ParseError("BRWT: children do not match the partition");  // BRWT.cpp:106-107
    if (lazy && depth == 0 && n_children == 0 && !node->raw.empty) {
        // a root that is a leaf (single column): its popcount sizes its own select samples, so decode it now
        decode_rrr63(node->raw, &node->bits);
        node->raw.clear();
    }
    node->children.resize(n_children);
    for (auto &c : node->children) load_brwt_node(in, &c, lazy, depth + 1);
}

}  // namespace ggq
