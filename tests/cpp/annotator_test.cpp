// Reference-side drop-in check at the CALLER level: the reference's own annotator class template
// instantiated over the GPU matrix — annotate::StaticBinRelAnnotator<ggq::GpuBRWT, std::string> — exactly as
// annotation/annotate_static.cpp:179-187 instantiates it over BRWT, with kFileExtension = ".brwt.annodbg"
// (annotation/static_annotators_def.hpp:15) and count_labels specialised to the device histogram.
//
// The template's member definitions live in the reference's annotate_static.cpp, so that TU is #included here
// (nothing is copied); everything else the annotators need is linked from the unmodified reference sources by
// oracle/Makefile (target annotator_test).  An annotation is built with the reference's ColumnCompressed +
// convert_to_simple_BRWT, written with the reference's serialize(), and loaded by the GPU annotator through
// merge_load() -> LabelEncoder::load + ggq::GpuBRWT::load on the same .brwt.annodbg file.  Expected answers are the
// known-answer cases of the reference's tests/test_annotation_BRWT.cpp (:195-218 serialization, :250-300 has_labels,
// :344-398 get_top_labels, :400-470 get_labels), each also compared with the reference annotator run here.
#include "annotate_static.cpp"   // the reference TU: template definitions (+ its own five instantiations)

#include <cstdio>
#include <set>

#include "annotate_column_compressed.hpp"
#include "annotation_converters.hpp"
#include "ggq/gpu_matrix.hpp"

namespace annotate {
typedef StaticBinRelAnnotator<ggq::GpuBRWT, std::string> GpuBRWTAnnotator;
template <>
const std::string GpuBRWTAnnotator::kFileExtension = kBRWTExtension;
// count_labels: one device histogram instead of one get_row per index (annotate_static.cpp:149-162)
template <>
std::vector<uint64_t> GpuBRWTAnnotator::count_labels(const std::vector<Index> &indices) const {
    return matrix_->count_columns(indices);
}
template class StaticBinRelAnnotator<ggq::GpuBRWT, std::string>;
}  // namespace annotate

using annotate::BRWTCompressed;
using annotate::GpuBRWTAnnotator;
typedef std::vector<std::string> Labels;
typedef std::vector<std::pair<std::string, size_t>> Counts;

static int g_fail = 0;
#define REQUIRE(c) do { if (!(c)) { std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); ++g_fail; } } while (0)

static std::set<std::string> as_set(const Labels &v) { return std::set<std::string>(v.begin(), v.end()); }

struct Pair {
    std::unique_ptr<BRWTCompressed<>> ref;
    std::unique_ptr<GpuBRWTAnnotator> gpu;
};

// rows[i] = labels of object i  ->  reference annotator + GPU annotator loaded from the reference's file
static Pair make(uint64_t n, const std::vector<std::pair<uint64_t, Labels>> &rows, const char *tag) {
    annotate::ColumnCompressed<> init(n);
    for (const auto &r : rows) init.add_labels(r.first, r.second);
    Pair p;
    p.ref = annotate::convert_to_simple_BRWT<BRWTCompressed<>>(std::move(init));
    const std::string base = std::string("/tmp/ggq_annotator_test_") + tag;
    p.ref->serialize(base);
    p.gpu.reset(new GpuBRWTAnnotator());
    REQUIRE(!p.gpu->load(base + "_missing"));
    REQUIRE(p.gpu->load(base));   // MultiLabelEncoded::load -> merge_load({base}) -> base + ".brwt.annodbg"
    REQUIRE(p.gpu->num_objects() == p.ref->num_objects());
    REQUIRE(p.gpu->num_labels() == p.ref->num_labels());
    REQUIRE(p.gpu->num_relations() == p.ref->num_relations());
    // serialize() through the annotator: same size as the reference's file, the matrix part (everything behind the
    // label encoder, whose hash-map order is not deterministic) byte for byte, and the reference annotator loads it
    p.gpu->serialize(base + "_again");
    std::ifstream a(base + ".brwt.annodbg", std::ios::binary), b(base + "_again.brwt.annodbg", std::ios::binary);
    std::string sa((std::istreambuf_iterator<char>(a)), std::istreambuf_iterator<char>());
    std::string sb((std::istreambuf_iterator<char>(b)), std::istreambuf_iterator<char>());
    const size_t mat = size_t(p.gpu->data().loaded_stream_bytes());
    REQUIRE(!sa.empty() && sa.size() == sb.size() && mat > 0 && mat <= sa.size());
    REQUIRE(sa.compare(sa.size() - mat, mat, sb, sb.size() - mat, mat) == 0);
    BRWTCompressed<> reloaded;
    REQUIRE(reloaded.load(base + "_again"));
    REQUIRE(reloaded.num_objects() == p.ref->num_objects() && reloaded.num_relations() == p.ref->num_relations());
    for (uint64_t i = 0; i < n; i += std::max<uint64_t>(1, n / 50)) REQUIRE(as_set(reloaded.get(i)) == as_set(p.ref->get(i)));
    std::remove((base + ".brwt.annodbg").c_str());
    std::remove((base + "_again.brwt.annodbg").c_str());
    return p;
}

static void serialization_and_get() {   // test_annotation_BRWT.cpp:195-218
    Pair p = make(5, {{0, {"Label0", "Label2", "Label8"}}, {2, {"Label1", "Label2"}}, {4, {"Label8"}}}, "ser");
    const std::vector<Labels> expect = {{"Label0", "Label2", "Label8"}, {}, {"Label1", "Label2"}, {}, {"Label8"}};
    for (uint64_t i = 0; i < 5; ++i) {
        REQUIRE(as_set(p.gpu->get(i)) == as_set(expect[i]));
        REQUIRE(as_set(p.gpu->get(i)) == as_set(p.ref->get(i)));
        REQUIRE(as_set(p.gpu->get_labels(i)) == as_set(expect[i]));
    }
    for (const char *l : {"Label0", "Label1", "Label2", "Label8", "Label9"})
        for (uint64_t i = 0; i < 5; ++i) REQUIRE(p.gpu->has_label(i, l) == p.ref->has_label(i, l));
    bool threw = false;
    try { p.gpu->add_label(0, "x"); } catch (const std::runtime_error &) { threw = true; }   // static representation
    REQUIRE(threw);
}

static void has_labels() {   // test_annotation_BRWT.cpp:250-300
    Pair p = make(5, {{0, {"Label0", "Label2", "Label8"}}, {2, {"Label1", "Label2"}}, {4, {"Label8"}}}, "has");
    struct Case { uint64_t i; Labels q; bool expect; };
    const Labels six = {"Label0", "Label1", "Label2", "Label4", "Label5", "Label8"}, four = {"Label0", "Label2", "Label4", "Label8"};
    const std::vector<Case> cases = {
        {0, six, false}, {0, four, false}, {0, {"Label0", "Label2", "Label8"}, true}, {0, {"Label0", "Label8"}, true},
        {0, {"Label2"}, true}, {0, {}, true},
        {1, six, false}, {1, four, false}, {1, {"Label0", "Label2", "Label8"}, false}, {1, {"Label0", "Label8"}, false},
        {1, {"Label2"}, false}, {1, {}, true},
        {2, six, false}, {2, four, false}, {2, {"Label1", "Label2", "Label8"}, false}, {2, {"Label1", "Label8"}, false},
        {2, {"Label1", "Label2"}, true}, {2, {"Label2"}, true}, {2, {}, true}};
    for (const Case &c : cases) {
        REQUIRE(p.gpu->has_labels(c.i, c.q) == c.expect);
        REQUIRE(p.ref->has_labels(c.i, c.q) == c.expect);
    }
}

static void top_labels() {   // test_annotation_BRWT.cpp:344-398
    Pair p = make(5, {{0, {"Label0", "Label2", "Label8"}}, {2, {"Label1", "Label2"}}, {3, {"Label1", "Label2", "Label8"}},
                      {4, {"Label2", "Label8"}}}, "top");
    const std::vector<uint64_t> all = {0, 1, 2, 3, 4};
    const Counts ranked = {{"Label2", 4}, {"Label8", 3}, {"Label1", 2}, {"Label0", 1}};
    REQUIRE(p.gpu->get_top_labels(all, 0) == Counts());
    REQUIRE(p.gpu->get_top_labels({}) == Counts());
    REQUIRE(p.gpu->get_top_labels(all) == ranked);
    for (size_t k : {1, 2, 3, 4, 1000})
        REQUIRE(p.gpu->get_top_labels(all, k) == Counts(ranked.begin(), ranked.begin() + std::min<size_t>(k, 4)));
    const Counts one = p.gpu->get_top_labels({2});
    REQUIRE((std::set<std::pair<std::string, size_t>>(one.begin(), one.end()) ==
             std::set<std::pair<std::string, size_t>>{{"Label1", 1}, {"Label2", 1}}));
    REQUIRE(p.gpu->get_top_labels(all) == p.ref->get_top_labels(all));   // (no ties among the counts here)
}

static void labels_by_ratio() {   // test_annotation_BRWT.cpp:400-470
    Pair p = make(5, {{0, {"Label0", "Label2", "Label8"}}, {2, {"Label1", "Label2"}}, {3, {"Label1", "Label2", "Label8"}},
                      {4, {"Label2"}}}, "ratio");
    struct Case { std::vector<uint64_t> rows; double ratio; Labels expect; };
    const std::vector<uint64_t> all = {0, 1, 2, 3, 4};
    const std::vector<Case> cases = {
        {{}, 1, {}},
        {{2}, 1, {"Label1", "Label2"}}, {{2}, 0, {"Label1", "Label2"}}, {{2}, 0.5, {"Label1", "Label2"}},
        {{2, 4}, 1, {"Label2"}}, {{2, 4}, 0, {"Label1", "Label2"}}, {{2, 4}, 0.5, {"Label1", "Label2"}},
        {{2, 4}, 0.501, {"Label2"}},
        {all, 1, {}}, {all, 0, {"Label0", "Label1", "Label2", "Label8"}}, {all, 0.2, {"Label0", "Label1", "Label2", "Label8"}},
        {all, 0.201, {"Label1", "Label2", "Label8"}}, {all, 0.4, {"Label1", "Label2", "Label8"}},
        {all, 0.401, {"Label2"}}, {all, 0.8, {"Label2"}}, {all, 0.801, {}}};
    for (const Case &c : cases) {
        REQUIRE(as_set(p.gpu->get_labels(c.rows, c.ratio)) == as_set(c.expect));
        REQUIRE(as_set(p.ref->get_labels(c.rows, c.ratio)) == as_set(c.expect));
    }
}

// beyond the reference's five-object cases: a random annotation, every call against the reference annotator
static void random_annotation() {
    const uint64_t n = 3000, m = 40;
    std::mt19937 gen(5);
    std::vector<std::pair<uint64_t, Labels>> rows;
    for (uint64_t i = 0; i < n; ++i) {
        Labels l;
        for (uint64_t j = 0; j < m; ++j)
            if (gen() % 100 < (j % 7 == 0 ? 30u : 4u)) l.push_back("L" + std::to_string(j));
        if (!l.empty()) rows.push_back({i, l});
    }
    Pair p = make(n, rows, "rand");
    for (uint64_t i = 0; i < n; i += 13) REQUIRE(as_set(p.gpu->get(i)) == as_set(p.ref->get(i)));
    std::vector<uint64_t> idx;
    for (int t = 0; t < 500; ++t) idx.push_back(gen() % n);
    for (double ratio : {0.0, 0.01, 0.05, 0.3, 1.0})
        REQUIRE(as_set(p.gpu->get_labels(idx, ratio)) == as_set(p.ref->get_labels(idx, ratio)));
    Counts a = p.gpu->get_top_labels(idx, 1000), b = p.ref->get_top_labels(idx, 1000);
    std::sort(a.begin(), a.end());   // tie order is unspecified in the reference (std::sort, annotate.cpp:69-72)
    std::sort(b.begin(), b.end());
    REQUIRE(a == b);
}

int main() {
    int n_dev = 0;
    if (ggq_device_count(&n_dev) != GGQ_OK || n_dev < 1) {
        std::fprintf(stderr, "annotator_test: no CUDA device (the engine has no CPU fallback)\n");
        return 2;
    }
    serialization_and_get();
    has_labels();
    top_labels();
    labels_by_ratio();
    random_annotation();
    if (g_fail) {
        std::fprintf(stderr, "annotator_test: %d check(s) FAILED\n", g_fail);
        return 1;
    }
    std::printf("annotator_test: all checks passed\n");
    return 0;
}
