// TEST INFRASTRUCTURE — the reference-side drop-in check, in C++, at the BinaryMatrix boundary.
//
// Compiled INSIDE the reference's include path (oracle/Makefile target `adapter_test`), linked against
// the unmodified reference objects and libggq.so.  For a grid of matrices it builds the reference BRWT /
// ColMajorCompressed, serializes it with the reference's own serialize(), loads the stream into
// ggq::GpuBRWT / ggq::GpuColMajor (include/ggq/gpu_matrix.hpp) through BinaryMatrix::load(std::istream&),
// and requires IDENTICAL answers through the BinaryMatrix interface — the same triple check as the
// reference's test_brwt() helper (tests/test_BRWT.cpp:92-150), but with vector equality (order included),
// plus serialize() byte equality and the batched extensions.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <random>
#include <sstream>

#include "BRWT.hpp"
#include "BRWT_builders.hpp"
#include "column_major.hpp"
#include "ggq/gpu_matrix.hpp"

static int g_fail = 0;
#define REQUIRE(cond)                                                                   \
    do {                                                                                \
        if (!(cond)) {                                                                  \
            std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);      \
            ++g_fail;                                                                   \
            return;                                                                     \
        }                                                                               \
    } while (0)

typedef std::vector<std::unique_ptr<bit_vector>> Columns;

static Columns make_columns(uint64_t n, uint64_t m, double d, unsigned seed) {
    std::mt19937 gen(seed);
    std::bernoulli_distribution bern(d);
    Columns cols;
    for (uint64_t j = 0; j < m; ++j) {
        sdsl::bit_vector bv(n, 0);
        for (uint64_t i = 0; i < n; ++i) bv[i] = bern(gen);
        cols.emplace_back(new bit_vector_stat(std::move(bv)));
    }
    return cols;
}

static void compare(const BinaryMatrix &ref, const ggq::GpuMatrixBase &gpu, const std::string &stream) {
    REQUIRE(ref.num_rows() == gpu.num_rows());
    REQUIRE(ref.num_columns() == gpu.num_columns());
    REQUIRE(ref.num_relations() == gpu.num_relations());
    const uint64_t n = ref.num_rows(), m = ref.num_columns();
    std::vector<uint64_t> rows(n);
    for (uint64_t i = 0; i < n; ++i) rows[i] = i;
    auto batched = gpu.get_rows(rows);
    auto sliced = gpu.slice_rows(rows);
    size_t sp = 0;
    for (uint64_t i = 0; i < n; ++i) {
        auto e = ref.get_row(i);
        REQUIRE(batched[i] == e);
        if (i % 17 == 0) REQUIRE(gpu.get_row(i) == e);
        for (auto c : e) REQUIRE(sliced[sp++] == c);
        REQUIRE(sliced[sp++] == ~uint64_t(0));
    }
    REQUIRE(sp == sliced.size());
    {   // the counter of StaticBinRelAnnotator::count_labels (annotate_static.cpp:149-162), duplicates counted
        std::vector<uint64_t> idx(rows);
        for (uint64_t i = 0; i < n; i += 3) idx.push_back(i);
        std::vector<uint64_t> expect(m, 0);
        for (auto i : idx)
            for (auto j : ref.get_row(i)) expect[j]++;
        REQUIRE(gpu.count_columns(idx) == expect);
        // get_labels(indices, ratio) threshold (annotate_static.cpp:79-91) and get_top_labels (annotate.cpp:57-83)
        for (double ratio : {0.0, 0.1, 0.5, 1.0}) {
            const size_t need = ratio == 0 ? 1 : size_t(std::ceil(idx.size() * ratio));
            std::vector<uint64_t> keep;
            for (uint64_t j = 0; j < m; ++j)
                if (expect[j] && expect[j] >= need) keep.push_back(j);
            REQUIRE(gpu.get_columns_with_ratio(idx, ratio) == keep);
        }
        std::vector<std::pair<uint64_t, uint64_t>> top;
        for (uint64_t j = 0; j < m; ++j)
            if (expect[j]) top.emplace_back(j, expect[j]);
        std::stable_sort(top.begin(), top.end(), [](const auto &a, const auto &b) { return a.second > b.second; });
        top.resize(std::min<size_t>(top.size(), 3));
        REQUIRE(gpu.get_top_columns(idx, 3) == top);
    }
    for (uint64_t j = 0; j < m; ++j) {
        REQUIRE(gpu.get_column(j) == ref.get_column(j));
        auto hc = gpu.has_column(rows, j);
        for (uint64_t i = 0; i < n; i += 7) {
            REQUIRE(hc[i] == ref.get(i, j));
            REQUIRE(gpu.get(i, j) == ref.get(i, j));
        }
    }
    std::ostringstream out;
    gpu.serialize(out);
    REQUIRE(out.str() == stream);
}

static void check_brwt(Columns &&cols, int arity, bool relax) {
    BRWT ref = BRWTBottomUpBuilder::build(std::move(cols), BRWTBottomUpBuilder::get_basic_partitioner(arity));
    if (relax) BRWTOptimizer::relax(&ref, 7);
    std::ostringstream out;
    ref.serialize(out);
    std::string stream = out.str();
    std::istringstream in(stream + "tail");
    ggq::GpuBRWT gpu;
    REQUIRE(gpu.load(in));
    REQUIRE(size_t(in.tellg()) == stream.size());   // load() consumed exactly the matrix
    REQUIRE(gpu.num_nodes() == ref.num_nodes());
    REQUIRE(gpu.total_column_size() == ref.total_column_size());
    compare(ref, gpu, stream);
    ggq::GpuBRWT copy = gpu;  // copies share the device image (BRWT is copyable, BRWT.hpp:27-31)
    REQUIRE(copy.get_column(0) == ref.get_column(0));
    // flattened device image next to the serialized matrix: restored without parsing, same answers, and
    // serialize() still reproduces the reference's stream byte for byte
    const std::string img = "/tmp/ggq_adapter_test.ggqimg", ser = "/tmp/ggq_adapter_test.brwt";
    gpu.save_image(img);
    { std::ofstream f(ser, std::ios::binary); f << stream; }
    ggq::GpuBRWT restored;
    REQUIRE(restored.load_image(img, ser));
    REQUIRE(restored.num_nodes() == ref.num_nodes());
    compare(ref, restored, stream);
    REQUIRE(!restored.load_image(ser));   // not an image
    // an image alone (no stream path): the image carries the host tree, serialize() re-encodes the device image
    ggq::GpuBRWT image_only;
    REQUIRE(image_only.load_image(img));
    REQUIRE(image_only.get_column(0) == ref.get_column(0));
    std::ostringstream sink;
    image_only.serialize(sink);
    REQUIRE(sink.str() == stream);
    // a column-major matrix without its stream has nothing to fall back to: serialize() throws, writes nothing
    {
        ggq::GpuColMajor cm;   // not loaded at all
        bool threw = false;
        std::ostringstream none;
        try { cm.serialize(none); } catch (const std::exception &) { threw = true; }
        REQUIRE(threw && none.str().empty());
    }
    // no host copy, no file: serialize() re-encodes the device image into the reference's stream format
    {
        ggq::GpuBRWT no_copy;
        no_copy.keep_stream(false);
        std::istringstream in2(stream);
        REQUIRE(no_copy.load(in2));
        std::ostringstream from_device;
        no_copy.serialize(from_device);
        REQUIRE(from_device.str() == stream);
        BRWT back;   // ... which the reference loads
        std::istringstream in3(from_device.str());
        REQUIRE(back.load(in3));
        REQUIRE(back.num_relations() == ref.num_relations());
    }
    // load_file(): no host copy of the stream; serialize() re-reads the file
    ggq::GpuBRWT from_file;
    REQUIRE(from_file.load_file(ser));
    REQUIRE(from_file.loaded_stream_bytes() == stream.size());
    std::ostringstream again;
    from_file.serialize(again);
    REQUIRE(again.str() == stream);
    std::remove(img.c_str());
    std::remove(ser.c_str());
}

static void check_colmajor(uint64_t n, uint64_t m, double d, unsigned seed) {
    std::mt19937 gen(seed);
    std::bernoulli_distribution bern(d);
    std::vector<std::unique_ptr<bit_vector_sd>> cols;
    for (uint64_t j = 0; j < m; ++j) {
        sdsl::bit_vector bv(n, 0);
        for (uint64_t i = 0; i < n; ++i) bv[i] = bern(gen);
        cols.emplace_back(new bit_vector_sd(bv));
    }
    ColMajorCompressed ref(std::move(cols));
    std::ostringstream out;
    ref.serialize(out);
    std::string stream = out.str();
    std::istringstream in(stream);
    ggq::GpuColMajor gpu;
    REQUIRE(gpu.load(in));
    compare(ref, gpu, stream);
}

static void check_bad_streams() {
    ggq::GpuBRWT gpu;
    std::istringstream empty("");
    REQUIRE(!gpu.load(empty));
    std::istringstream junk(std::string(64, '\xff'));
    REQUIRE(!gpu.load(junk));
    bool threw = false;
    try { gpu.get_row(0); } catch (const std::runtime_error &) { threw = true; }
    REQUIRE(threw);
}

int main() {
    int n_dev = 0;
    if (ggq_device_count(&n_dev) != GGQ_OK || n_dev < 1) {
        std::fprintf(stderr, "adapter_test: no CUDA device (the engine has no CPU fallback)\n");
        return 2;
    }
    for (uint64_t m : {1, 2, 5, 19, 64, 100})
        for (uint64_t n : {1, 19, 500})
            for (double d : {0.0, 0.1, 1.0})
                check_brwt(make_columns(n, m, d, unsigned(n * 131 + m)), m % 3 + 2, m > 5);
    check_brwt(make_columns(3000, 300, 0.02, 7), 4, true);
    check_colmajor(500, 20, 0.1, 1);
    check_colmajor(500, 20, 0.9, 2);
    check_colmajor(64, 3, 0.0, 3);
    check_bad_streams();
    if (g_fail) {
        std::fprintf(stderr, "adapter_test: %d check(s) FAILED\n", g_fail);
        return 1;
    }
    std::printf("adapter_test: all checks passed\n");
    return 0;
}
