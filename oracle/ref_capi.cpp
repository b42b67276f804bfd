// TEST INFRASTRUCTURE — not part of the product.
//
// C-ABI glue over the UNMODIFIED reference sources (compiled from
// /root/reference by oracle/Makefile into oracle/_ref/libggref.so).  It lets
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// leg drive the reference's own BRWT / ColMajorCompressed / bit_vector_* code
// through ctypes.  Nothing under genome_graph_annotation_b200/ may link or load it.
//
// Reference interfaces driven here:
//   BinaryMatrix                         common/binary_matrix.hpp:9-29
//   BRWT                                 annotation/hierarchical_annotation/BRWT.hpp:18-61
//   BRWTBottomUpBuilder / BRWTOptimizer  annotation/hierarchical_annotation/BRWT_builders.hpp:38-83
//   binary_grouping_greedy               annotation/hierarchical_annotation/partitionings.hpp:15-16
//   ColMajorCompressed                   annotation/column_compressed/column_major.hpp:10-37
//   bit_vector_{stat,sd,rrr<63>}         common/bit_vector.hpp:83-225
//   DataGenerator                        experiments/data_generation.hpp:7-72
//   experiments' generate_brwt_from_rows experiments/method_constructors.hpp:110-133

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include <chrono>
#include <omp.h>

#include "binary_matrix.hpp"
#include "bit_vector.hpp"
#include "BRWT.hpp"
#include "BRWT_builders.hpp"
#include "partitionings.hpp"
#include "column_major.hpp"
#include "data_generation.hpp"
#include "serialization.hpp"

namespace {

sdsl::bit_vector words_to_sdsl(const uint64_t *words, uint64_t n_bits) {
    sdsl::bit_vector bv(n_bits, 0);
    if (n_bits)
        std::memcpy(bv.data(), words, ((n_bits + 63) / 64) * 8);
    // clear the padding bits of the last word so that popcounts are exact
    if (n_bits % 64)
        bv.data()[(n_bits - 1) / 64] &= (~uint64_t(0)) >> (64 - n_bits % 64);
    return bv;
}

std::vector<std::unique_ptr<bit_vector>>
columns_from_words(const uint64_t *col_words, uint64_t n, uint64_t m) {
    const uint64_t wpc = (n + 63) / 64;
    std::vector<std::unique_ptr<bit_vector>> columns(m);
    for (uint64_t j = 0; j < m; ++j)
        columns[j] = std::make_unique<bit_vector_stat>(words_to_sdsl(col_words + j * wpc, n));
    return columns;
}

void columns_to_words(const std::vector<std::unique_ptr<bit_vector>> &cols,
                      uint64_t n, uint64_t *out_words) {
    const uint64_t wpc = (n + 63) / 64;
    std::memset(out_words, 0, wpc * cols.size() * 8);
    for (uint64_t j = 0; j < cols.size(); ++j) {
        uint64_t *dst = out_words + j * wpc;
        cols[j]->call_ones([&](uint64_t i) { dst[i >> 6] |= uint64_t(1) << (i & 63); });
    }
}

uint8_t *dup_bytes(const std::string &s, uint64_t *len) {
    uint8_t *p = static_cast<uint8_t *>(std::malloc(s.size() ? s.size() : 1));
    std::memcpy(p, s.data(), s.size());
    *len = s.size();
    return p;
}

struct BV {
    int kind;  // 0 stat, 1 sd, 2 rrr
    std::unique_ptr<bit_vector> v;
};

}  // namespace

extern "C" {

// ---------------------------------------------------------------- matrices --

void *ggref_brwt_build(const uint64_t *col_words, uint64_t n, uint64_t m,
                       int arity, int greedy, uint64_t relax_arity, int threads) {
    auto columns = columns_from_words(col_words, n, m);
    std::unique_ptr<BRWT> mat;
    if (greedy) {
        // experiments/method_constructors.hpp:115-120 uses binary_grouping_greedy
        // (single-threaded); the parallel variant is the same partition function.
        mat = std::make_unique<BRWT>(BRWTBottomUpBuilder::build(
            std::move(columns),
            threads > 1 ? get_parallel_binary_grouping_greedy(threads)
                        : BRWTBottomUpBuilder::Partitioner(binary_grouping_greedy),
            threads > 1 ? threads : 1));
    } else {
        mat = std::make_unique<BRWT>(BRWTBottomUpBuilder::build(
            std::move(columns), BRWTBottomUpBuilder::get_basic_partitioner(arity),
            threads > 1 ? threads : 1));
    }
    if (relax_arity > 1)
        BRWTOptimizer::relax(mat.get(), relax_arity, threads > 1 ? threads : 1);
    return static_cast<BinaryMatrix *>(mat.release());
}

void *ggref_colmajor_build(const uint64_t *col_words, uint64_t n, uint64_t m) {
    const uint64_t wpc = (n + 63) / 64;
    std::vector<std::unique_ptr<bit_vector_sd>> cols(m);
    for (uint64_t j = 0; j < m; ++j)
        cols[j] = std::make_unique<bit_vector_sd>(words_to_sdsl(col_words + j * wpc, n));
    return static_cast<BinaryMatrix *>(new ColMajorCompressed(std::move(cols)));
}

// kind: 0 = BRWT, 1 = ColMajorCompressed.  Returns NULL when load() returns false.
void *ggref_matrix_load(int kind, const uint8_t *buf, uint64_t len) {
    std::unique_ptr<BinaryMatrix> mat;
    if (kind == 0) mat.reset(new BRWT());
    else mat.reset(new ColMajorCompressed());
    std::istringstream in(std::string(reinterpret_cast<const char *>(buf), len));
    if (!mat->load(in))
        return nullptr;
    return mat.release();
}

uint8_t *ggref_matrix_serialize(const void *h, uint64_t *len) {
    std::ostringstream out;
    static_cast<const BinaryMatrix *>(h)->serialize(out);
    return dup_bytes(out.str(), len);
}

void ggref_matrix_free(void *h) { delete static_cast<BinaryMatrix *>(h); }
void ggref_bytes_free(void *p) { std::free(p); }

uint64_t ggref_matrix_num_rows(const void *h) { return static_cast<const BinaryMatrix *>(h)->num_rows(); }
uint64_t ggref_matrix_num_columns(const void *h) { return static_cast<const BinaryMatrix *>(h)->num_columns(); }
uint64_t ggref_matrix_num_relations(const void *h) { return static_cast<const BinaryMatrix *>(h)->num_relations(); }

// BRWT-only stats (BRWT.hpp:44-49); return 0 for other matrix types.
uint64_t ggref_brwt_num_nodes(const void *h) {
    auto *b = dynamic_cast<const BRWT *>(static_cast<const BinaryMatrix *>(h));
    return b ? b->num_nodes() : 0;
}
uint64_t ggref_brwt_total_column_size(const void *h) {
    auto *b = dynamic_cast<const BRWT *>(static_cast<const BinaryMatrix *>(h));
    return b ? b->total_column_size() : 0;
}
double ggref_brwt_avg_arity(const void *h) {
    auto *b = dynamic_cast<const BRWT *>(static_cast<const BinaryMatrix *>(h));
    return b ? b->avg_arity() : 0;
}

int ggref_matrix_get(const void *h, uint64_t row, uint64_t col) {
    return static_cast<const BinaryMatrix *>(h)->get(row, col) ? 1 : 0;
}

// out may be NULL (count only).  Returns the number of set columns.
uint64_t ggref_matrix_get_row(const void *h, uint64_t row, uint64_t *out, uint64_t cap) {
    auto r = static_cast<const BinaryMatrix *>(h)->get_row(row);
    if (out)
        std::memcpy(out, r.data(), std::min<uint64_t>(cap, r.size()) * 8);
    return r.size();
}

// malloc'ed result; free with ggref_bytes_free.
uint64_t *ggref_matrix_get_column(const void *h, uint64_t col, uint64_t *n_out) {
    auto r = static_cast<const BinaryMatrix *>(h)->get_column(col);
    *n_out = r.size();
    uint64_t *p = static_cast<uint64_t *>(std::malloc(r.size() ? r.size() * 8 : 8));
    std::memcpy(p, r.data(), r.size() * 8);
    return p;
}

// The "reference OpenMP CPU path" of BASELINE.md §3: omp parallel for around
// the UNMODIFIED scalar get_row (BRWT.cpp:26-53 / column_major.cpp:34-47).
// offsets has n+1 entries; *cols_out is malloc'ed (ggref_bytes_free).
// Returns the wall-clock seconds of the query loop only.
double ggref_matrix_get_rows(const void *h, const uint64_t *rows, uint64_t n,
                             uint64_t *offsets, uint64_t **cols_out, int threads) {
    const BinaryMatrix *mat = static_cast<const BinaryMatrix *>(h);
    std::vector<std::vector<uint64_t>> res(n);
    auto t0 = std::chrono::steady_clock::now();
    #pragma omp parallel for num_threads(threads) schedule(dynamic, 4096)
    for (uint64_t t = 0; t < n; ++t)
        res[t] = mat->get_row(rows[t]);
    auto t1 = std::chrono::steady_clock::now();
    offsets[0] = 0;
    for (uint64_t t = 0; t < n; ++t)
        offsets[t + 1] = offsets[t] + res[t].size();
    if (cols_out) {
        uint64_t *p = static_cast<uint64_t *>(std::malloc(offsets[n] ? offsets[n] * 8 : 8));
        for (uint64_t t = 0; t < n; ++t)
            std::memcpy(p + offsets[t], res[t].data(), res[t].size() * 8);
        *cols_out = p;
    }
    return std::chrono::duration<double>(t1 - t0).count();
}

// Timing-only variant: consumes the rows into a checksum, keeps nothing.
double ggref_matrix_time_get_rows(const void *h, const uint64_t *rows, uint64_t n,
                                  int threads, uint64_t *checksum) {
    const BinaryMatrix *mat = static_cast<const BinaryMatrix *>(h);
    uint64_t sum = 0;
    auto t0 = std::chrono::steady_clock::now();
    #pragma omp parallel for num_threads(threads) schedule(dynamic, 4096) reduction(+:sum)
    for (uint64_t t = 0; t < n; ++t) {
        auto r = mat->get_row(rows[t]);
        uint64_t s = r.size();
        for (auto c : r) s += c;
        sum += s;
    }
    auto t1 = std::chrono::steady_clock::now();
    if (checksum) *checksum = sum;
    return std::chrono::duration<double>(t1 - t0).count();
}

double ggref_matrix_time_get(const void *h, const uint64_t *rows, const uint64_t *cols,
                             uint64_t n, int threads, uint8_t *out) {
    const BinaryMatrix *mat = static_cast<const BinaryMatrix *>(h);
    auto t0 = std::chrono::steady_clock::now();
    #pragma omp parallel for num_threads(threads) schedule(dynamic, 4096)
    for (uint64_t t = 0; t < n; ++t)
        out[t] = mat->get(rows[t], cols[t]);
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// ------------------------------------------------------------- bit vectors --

void *ggref_bv_build(int kind, const uint64_t *words, uint64_t n_bits) {
    auto *b = new BV{kind, nullptr};
    sdsl::bit_vector bv = words_to_sdsl(words, n_bits);
    if (kind == 0) b->v.reset(new bit_vector_stat(std::move(bv)));
    else if (kind == 1) b->v.reset(new bit_vector_sd(bv));
    else b->v.reset(new bit_vector_rrr<>(bv));
    return b;
}

void *ggref_bv_load(int kind, const uint8_t *buf, uint64_t len) {
    auto b = std::make_unique<BV>();
    b->kind = kind;
    if (kind == 0) b->v.reset(new bit_vector_stat());
    else if (kind == 1) b->v.reset(new bit_vector_sd());
    else b->v.reset(new bit_vector_rrr<>());
    std::istringstream in(std::string(reinterpret_cast<const char *>(buf), len));
    try {
        if (!b->v->load(in)) return nullptr;
    } catch (...) {
        return nullptr;
    }
    return b.release();
}

uint8_t *ggref_bv_serialize(const void *h, uint64_t *len) {
    std::ostringstream out;
    static_cast<const BV *>(h)->v->serialize(out);
    return dup_bytes(out.str(), len);
}

void ggref_bv_free(void *h) { delete static_cast<BV *>(h); }
uint64_t ggref_bv_size(const void *h) { return static_cast<const BV *>(h)->v->size(); }
uint64_t ggref_bv_num_set_bits(const void *h) { return static_cast<const BV *>(h)->v->num_set_bits(); }
int ggref_bv_is_inverted(const void *h) {
    auto *sd = dynamic_cast<const bit_vector_sd *>(static_cast<const BV *>(h)->v.get());
    return sd ? sd->is_inverted() : 0;
}

// The rest of the bit_vector interface (common/bit_vector.hpp:18-44,199), for the parity tests of ggq_bv_rank0 /
// select0 / get_int / call_ones_into.  op: 3 = rank0, 4 = select0 (bit_vector_rrr only, else returns -1),
// 5 = get_int(q[i], width).
int ggref_bv_query_aux(const void *h, int op, const uint64_t *q, uint64_t n, uint32_t width, uint64_t *out) {
    const bit_vector &v = *static_cast<const BV *>(h)->v;
    if (op == 3) {
        for (uint64_t i = 0; i < n; ++i) out[i] = v.rank0(q[i]);
    } else if (op == 4) {
        auto *rrr = dynamic_cast<const bit_vector_rrr<> *>(&v);
        if (!rrr) return -1;
        for (uint64_t i = 0; i < n; ++i) out[i] = rrr->select0(q[i]);
    } else if (op == 5) {
        for (uint64_t i = 0; i < n; ++i) out[i] = v.get_int(q[i], width);
    } else {
        return -1;
    }
    return 0;
}

// call_ones (bit_vector.hpp:44): positions in callback order; out must hold num_set_bits entries.
uint64_t ggref_bv_call_ones(const void *h, uint64_t *out) {
    uint64_t k = 0;
    static_cast<const BV *>(h)->v->call_ones([&](uint64_t i) { out[k++] = i; });
    return k;
}

// op: 0 = rank1, 1 = select1, 2 = operator[].  Returns seconds of the loop.
double ggref_bv_query(const void *h, int op, const uint64_t *q, uint64_t n,
                      uint64_t *out, int threads) {
    const bit_vector &v = *static_cast<const BV *>(h)->v;
    if (op == 0) v.rank1(0);  // force bit_vector_stat's lazy init outside the loop
    auto t0 = std::chrono::steady_clock::now();
    if (op == 0) {
        #pragma omp parallel for num_threads(threads) schedule(static)
        for (uint64_t i = 0; i < n; ++i) out[i] = v.rank1(q[i]);
    } else if (op == 1) {
        #pragma omp parallel for num_threads(threads) schedule(static)
        for (uint64_t i = 0; i < n; ++i) out[i] = v.select1(q[i]);
    } else {
        #pragma omp parallel for num_threads(threads) schedule(static)
        for (uint64_t i = 0; i < n; ++i) out[i] = v[q[i]];
    }
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// --------------------------------------------------------- data generators --

// experiments/main.cpp:224-231 ("norepl"): column-major words, wpc = ceil(n/64).
void ggref_gen_norepl(uint64_t n, uint64_t m, double d, int seed, uint64_t *out_words) {
    DataGenerator g;
    g.set_seed(seed);
    auto cols = g.generate_random_columns(n, m, std::vector<double>(m, d));
    columns_to_words(cols, n, out_words);
}

// experiments/main.cpp:232-247 ("uniform_rows"): n_unique distinct rows, each n/n_unique times.
void ggref_gen_uniform_rows(uint64_t n, uint64_t m, double d, uint64_t n_unique, int seed,
                            uint64_t *out_words) {
    DataGenerator g;
    g.set_seed(seed);
    auto cols = g.generate_random_rows(n_unique, m, std::vector<double>(m, d),
                                       std::vector<uint32_t>(n_unique, n / n_unique));
    columns_to_words(cols, n, out_words);
}

// experiments/main.cpp:78-83: positions for the sampled row/point/column queries.
void ggref_gen_random_ints(uint64_t count, uint64_t begin, uint64_t end, int seed, uint64_t *out) {
    DataGenerator g;
    g.set_seed(seed);
    auto v = g.generate_random_ints(count, begin, end);
    std::memcpy(out, v.data(), count * 8);
}

// Portable implicit matrix of SURVEY.md §8(d) (configs C4/C5): cell (i, j) is set iff
// mix64(seed * 0x9E3779B97F4A7C15 + i * m + j) < threshold.  Column-major words, wpc = ceil(n/64).
static inline uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL;
    x ^= x >> 27; x *= 0x94d049bb133111ebULL;
    x ^= x >> 31;
    return x;
}
void ggref_gen_implicit(uint64_t n, uint64_t m, uint64_t threshold, uint64_t seed, uint64_t *out_words) {
    const uint64_t wpc = (n + 63) / 64;
    const uint64_t base = seed * 0x9E3779B97F4A7C15ULL;
    #pragma omp parallel for schedule(static)
    for (uint64_t j = 0; j < m; ++j) {
        uint64_t *dst = out_words + j * wpc;
        for (uint64_t w = 0; w < wpc; ++w) {
            uint64_t word = 0;
            const uint64_t i0 = w * 64, i1 = std::min<uint64_t>(n, i0 + 64);
            for (uint64_t i = i0; i < i1; ++i)
                word |= uint64_t(mix64(base + i * m + j) < threshold) << (i - i0);
            dst[w] = word;
        }
    }
}

int ggref_max_threads() { return omp_get_max_threads(); }
// BRWTOptimizer's space model adds sizeof(bit_vector_rrr<>) * 8 bits per vector (BRWT_builders.cpp:293-300); the device
// builder restates the model and carries the number as a constant — this pins it.
uint64_t ggref_sizeof_bit_vector_rrr() { return sizeof(bit_vector_rrr<>); }

}  // extern "C"
