// ggq::LabelBatcher (include/ggq/label_batcher.hpp) under the reference's calling pattern: T pool threads, each
// answering one query sequence at a time (main.cpp:462-504 -> AnnotatedDBG::get_labels, common/annotated_dbg.cpp:88-116).
// Checks that every caller gets exactly the answer of its own ggq_get_labels call, and reports the sustained
// indices/s with and without the batcher.  Links only libggq.so (no reference code).
//   batcher_test [rows] [cols] [density] [threads] [queries per thread] [indices per query]
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <thread>

#include "ggq/label_batcher.hpp"

struct Query {
    std::vector<uint64_t> idx;
    double ratio;
    std::vector<uint64_t> cols, counts;   // expected
};

static void die(const char *what) {
    std::fprintf(stderr, "batcher_test: %s: %s\n", what, ggq_last_error());
    std::exit(1);
}

int main(int argc, char **argv) {
    const uint64_t n = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : 2000000;
    const uint64_t m = argc > 2 ? std::strtoull(argv[2], nullptr, 10) : 500;
    const double d = argc > 3 ? std::atof(argv[3]) : 0.01;
    const int T = argc > 4 ? std::atoi(argv[4]) : 8;
    const int Q = argc > 5 ? std::atoi(argv[5]) : 200;
    const int L = argc > 6 ? std::atoi(argv[6]) : 300;
    int n_dev = 0;
    if (ggq_device_count(&n_dev) != GGQ_OK || n_dev < 1) {
        std::fprintf(stderr, "batcher_test: no CUDA device (the engine has no CPU fallback)\n");
        return 2;
    }
    ggq_matrix *mat = nullptr;
    const uint64_t thr = uint64_t(std::ldexp((long double)d, 64));
    if (ggq_brwt_synthesize(n, m, thr, 42, 4, 0, &mat) != GGQ_OK) die("synthesize");

    // queries: L consecutive-ish row ids (the k-mers of one read), a few presence ratios
    std::vector<std::vector<Query>> work(T, std::vector<Query>(Q));
    std::mt19937_64 gen(7);
    const double ratios[4] = {0.0, 0.05, 0.25, 0.8};
    std::vector<uint64_t> cbuf(m), vbuf(m);
    for (int t = 0; t < T; ++t)
        for (int q = 0; q < Q; ++q) {
            Query &w = work[t][q];
            const int len = (q % 17 == 0) ? 1 + int(gen() % 40) : L;
            uint64_t r = gen() % n;
            for (int i = 0; i < len; ++i) { w.idx.push_back(r); r = (r + 1 + gen() % 3) % n; }
            w.ratio = ratios[gen() % 4];
            uint64_t k = 0;
            if (ggq_get_labels(mat, w.idx.data(), w.idx.size(), w.ratio, cbuf.data(), vbuf.data(), &k) != GGQ_OK) die("get_labels");
            w.cols.assign(cbuf.begin(), cbuf.begin() + k);
            w.counts.assign(vbuf.begin(), vbuf.begin() + k);
        }
    uint64_t total_idx = 0;
    for (auto &tw : work) for (auto &w : tw) total_idx += w.idx.size();

    std::atomic<int> wrong{0};
    auto timed = [&](bool batched, ggq::LabelBatcher *lb) {
        std::vector<std::thread> th;
        const auto t0 = std::chrono::steady_clock::now();
        for (int t = 0; t < T; ++t)
            th.emplace_back([&, t] {
                std::vector<uint64_t> c, v, cb(m), vb(m);
                for (int q = 0; q < Q; ++q) {
                    const Query &w = work[t][q];
                    if (batched) {
                        try {
                            lb->get_labels(w.idx, w.ratio, &c, &v);
                        } catch (const std::exception &e) {
                            std::fprintf(stderr, "batcher_test: thread %d query %d: %s\n", t, q, e.what());
                            ++wrong;
                            continue;
                        }
                    } else {
                        uint64_t k = 0;
                        if (ggq_get_labels(mat, w.idx.data(), w.idx.size(), w.ratio, cb.data(), vb.data(), &k) != GGQ_OK) {
                            std::fprintf(stderr, "batcher_test: thread %d query %d: %s\n", t, q, ggq_last_error());
                            ++wrong;
                            continue;
                        }
                        c.assign(cb.begin(), cb.begin() + k);
                        v.assign(vb.begin(), vb.begin() + k);
                    }
                    if (c != w.cols || v != w.counts) ++wrong;
                }
            });
        for (auto &x : th) x.join();
        return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    };
    timed(false, nullptr);   // warm-up: workspaces of every thread grow once
    const double s_direct = timed(false, nullptr);
    double s_batched = 0;
    ggq::LabelBatcher::Stats st;
    {
        ggq::LabelBatcher lb(mat, 64, std::chrono::microseconds(50));
        timed(true, &lb);
        ggq::LabelBatcher lb2(mat, 64, std::chrono::microseconds(50));
        s_batched = timed(true, &lb2);
        st = lb2.stats();
    }
    std::printf("{\"threads\": %d, \"queries\": %d, \"indices\": %llu, \"matrix\": \"%llu x %llu d=%g\", "
                "\"per_thread_calls_indices_per_s\": %.4g, \"batcher_indices_per_s\": %.4g, \"batches\": %llu, "
                "\"queries_per_batch\": %.2f, \"wrong\": %d}\n",
                T, T * Q, (unsigned long long)total_idx, (unsigned long long)n, (unsigned long long)m, d,
                total_idx / s_direct, total_idx / s_batched, (unsigned long long)st.batches,
                st.batches ? double(st.queries) / st.batches : 0.0, wrong.load());
    ggq_free(mat);
    if (wrong.load()) { std::fprintf(stderr, "batcher_test: %d wrong answers\n", wrong.load()); return 1; }
    std::printf("batcher_test: all checks passed\n");
    return 0;
}
