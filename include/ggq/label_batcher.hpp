// Cross-query batching for the per-sequence callers.
//
// `annograph classify` runs one task per query sequence on a thread pool (main.cpp:462-504); each task calls
// AnnotatedDBG::get_labels(sequence, ratio) -> annotator->get_labels(indices, ratio') with the 10^2..10^3 row indices
// of ONE sequence (common/annotated_dbg.cpp:88-116).  A batch that small cannot fill a B200, and every call pays a
// launch + synchronisation round trip.  LabelBatcher sits between those concurrent callers and the engine:
//
//   * a caller deposits its indices and blocks;
//   * the first caller of a batch is its leader: it waits until `max_queries` are collected or `max_wait` has passed
//     since it arrived, then submits the whole batch through ONE ggq_get_labels_batch_ratios call (one descent
//     launch, one histogram CTA per query) and hands every caller its own answer;
//   * while a batch is on the device, later callers already form the next one (the engine's workspace pool gives
//     every in-flight batch its own stream), so the device is never idle behind a host round trip.
//
// The answer of a query is exactly ggq_get_labels(indices, ratio): kept columns ascending + their counts
// (StaticBinRelAnnotator::get_labels, annotate_static.cpp:71-94).  No thread is created: batches are driven by the
// callers themselves.  Thread-safe; the matrix handle must outlive the batcher.
#ifndef GGQ_LABEL_BATCHER_HPP_
#define GGQ_LABEL_BATCHER_HPP_

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../ggq.h"

namespace ggq {

class LabelBatcher {
  public:
    struct Stats {
        uint64_t queries = 0, batches = 0, indices = 0;
    };

    // max_wait: longest a leader holds its batch open; idle_gap: it also submits as soon as nobody has joined for
    // this long (with T blocking callers a batch can never grow beyond T: waiting out max_wait would be pure latency)
    explicit LabelBatcher(const ggq_matrix *matrix, size_t max_queries = 64,
                          std::chrono::microseconds max_wait = std::chrono::microseconds(50),
                          std::chrono::microseconds idle_gap = std::chrono::microseconds(4))
        : m_(matrix), max_queries_(max_queries ? max_queries : 1), max_wait_(max_wait), idle_gap_(idle_gap) {
        if (!matrix) throw std::invalid_argument("ggq::LabelBatcher: null matrix");
    }
    LabelBatcher(const LabelBatcher &) = delete;
    LabelBatcher &operator=(const LabelBatcher &) = delete;

    // Blocking.  cols / counts receive the kept columns (ascending) and their counts; counts may be null.
    void get_labels(const uint64_t *indices, size_t n, double presence_ratio, std::vector<uint64_t> *cols,
                    std::vector<uint64_t> *counts = nullptr);

    void get_labels(const std::vector<uint64_t> &indices, double presence_ratio, std::vector<uint64_t> *cols,
                    std::vector<uint64_t> *counts = nullptr) {
        get_labels(indices.data(), indices.size(), presence_ratio, cols, counts);
    }

    Stats stats() const {
        std::lock_guard<std::mutex> lk(mu_);
        return stats_;
    }

  private:
    struct Batch {
        std::vector<uint64_t> rows, offsets{0};
        std::vector<double> ratios;
        // results
        std::vector<uint64_t> label_offsets, cols, counts;
        std::atomic<bool> closed{false}, done{false};
        std::atomic<uint32_t> joined{0};
        std::string error;
        std::condition_variable done_cv;
    };

    void run(Batch &b);

    const ggq_matrix *m_;
    const size_t max_queries_;
    const std::chrono::microseconds max_wait_, idle_gap_;
    mutable std::mutex mu_;
    std::shared_ptr<Batch> open_;   // the batch callers are currently joining (null: the next caller leads a new one)
    Stats stats_;
};

inline void LabelBatcher::run(Batch &b) {
    const uint64_t nq = b.ratios.size();
    b.label_offsets.assign(nq + 1, 0);
    uint64_t cap = std::max<uint64_t>(nq * std::min<uint64_t>(ggq_num_columns(m_), 64), 1024), kept = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        b.cols.resize(cap);
        b.counts.resize(cap);
        const int rc = ggq_get_labels_batch_ratios(m_, b.rows.data(), b.offsets.data(), nq, b.ratios.data(),
                                                   b.label_offsets.data(), b.cols.data(), b.counts.data(), cap, &kept);
        if (rc == GGQ_OK) return;
        if (rc == GGQ_ERR_ARG && kept > cap) { cap = kept; continue; }   // capacity too small: *n_out is what it needs
        b.error = std::string("ggq: ") + ggq_last_error();
        return;
    }
    b.error = "ggq: label batch did not fit its own size";
}

inline void LabelBatcher::get_labels(const uint64_t *indices, size_t n, double presence_ratio,
                                     std::vector<uint64_t> *cols, std::vector<uint64_t> *counts) {
    if (!cols) throw std::invalid_argument("ggq::LabelBatcher::get_labels: null result");
    std::unique_lock<std::mutex> lk(mu_);
    std::shared_ptr<Batch> b = open_;
    const bool leader = !b;
    if (leader) {
        b = std::make_shared<Batch>();
        open_ = b;
    }
    const size_t slot = b->ratios.size();
    b->rows.insert(b->rows.end(), indices, indices + n);
    b->offsets.push_back(b->rows.size());
    b->ratios.push_back(presence_ratio);
    b->joined.fetch_add(1, std::memory_order_release);
    if (b->ratios.size() >= max_queries_) {   // full: nobody else may join; the leader sees the flag
        b->closed.store(true, std::memory_order_release);
        if (open_ == b) open_.reset();
    }
    if (leader) {
        // The window is tens of microseconds — shorter than a futex wake-up — so the leader polls the flag instead of
        // sleeping on a condition variable.
        const auto deadline = std::chrono::steady_clock::now() + max_wait_;
        lk.unlock();
        uint32_t seen = 1;
        auto last_join = std::chrono::steady_clock::now();
        while (!b->closed.load(std::memory_order_acquire)) {
            const auto now = std::chrono::steady_clock::now();
            const uint32_t j = b->joined.load(std::memory_order_acquire);
            if (j != seen) { seen = j; last_join = now; }
            if (now >= deadline || now - last_join >= idle_gap_) break;
            std::this_thread::yield();
        }
        lk.lock();
        b->closed.store(true, std::memory_order_release);
        if (open_ == b) open_.reset();
        stats_.batches += 1;
        stats_.queries += b->ratios.size();
        stats_.indices += b->rows.size();
        lk.unlock();
        run(*b);   // the device call: no lock held, later callers form the next batch meanwhile
        lk.lock();
        b->done.store(true, std::memory_order_release);
        b->done_cv.notify_all();
    } else {
        // followers: poll for a while (a batch takes ~50-100 us), then sleep
        lk.unlock();
        const auto spin_until = std::chrono::steady_clock::now() + std::chrono::microseconds(500);
        while (!b->done.load(std::memory_order_acquire) && std::chrono::steady_clock::now() < spin_until)
            std::this_thread::yield();
        lk.lock();
        while (!b->done.load(std::memory_order_acquire)) b->done_cv.wait(lk);
    }
    if (!b->error.empty()) throw std::runtime_error(b->error);
    const uint64_t lo = b->label_offsets[slot], hi = b->label_offsets[slot + 1];
    cols->assign(b->cols.begin() + lo, b->cols.begin() + hi);
    if (counts) counts->assign(b->counts.begin() + lo, b->counts.begin() + hi);
}

}  // namespace ggq

#endif  // GGQ_LABEL_BATCHER_HPP_
