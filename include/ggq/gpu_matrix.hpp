// Host C++17 adapters: GPU-resident drop-ins for the reference's BinaryMatrix implementations.
//
//   GpuBRWT      replaces  BRWT                 (annotation/hierarchical_annotation/BRWT.hpp:18-61)
//   GpuColMajor  replaces  ColMajorCompressed   (annotation/column_compressed/column_major.hpp:10-37)
//
// Both derive from the reference's own `BinaryMatrix` (common/binary_matrix.hpp:9-29) — this header is
// compiled INSIDE the reference tree (it includes "binary_matrix.hpp" from the reference's include path)
// and links against libggq.so.  Same names, argument meaning and error behaviour as the reference:
// load() returns false on any failure and never throws (BRWT.cpp:88-110); serialize() throws
// std::ofstream::failure on a bad stream (BRWT.cpp:114-115); results are returned by value; const
// queries are safe from concurrent threads (the engine keeps a per-call stream + scratch pool).
// Batched calls named by the north star and absent in the reference are defined by composition:
//   get_rows(R)[t] == get_row(R[t]);  slice_rows(R) == concat_t(get_row(R[t]) ++ [UINT64_MAX]);
//   has_column(R, j)[t] == get(R[t], j).
// Runtime CUDA failures surface as std::runtime_error (there is no CPU fallback to hide them).
#ifndef GGQ_GPU_MATRIX_HPP_
#define GGQ_GPU_MATRIX_HPP_

#include <algorithm>
#include <fstream>
#include <istream>
#include <memory>
#include <ostream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "binary_matrix.hpp"  // the reference's interface (common/binary_matrix.hpp)
#include "../ggq.h"

namespace ggq {

class GpuMatrixBase : public BinaryMatrix {
  public:
    explicit GpuMatrixBase(int device = 0) : device_(device) {}

    uint64_t num_columns() const override { return h_ ? ggq_num_columns(h_.get()) : 0; }
    uint64_t num_rows() const override { return h_ ? ggq_num_rows(h_.get()) : 0; }
    uint64_t num_relations() const override { return h_ ? ggq_num_relations(h_.get()) : 0; }

    bool get(Row row, Column column) const override {
        uint8_t out = 0;
        check(ggq_get(handle(), &row, &column, 1, &out));
        return out;
    }

    std::vector<Column> get_row(Row row) const override {
        std::vector<std::vector<Column>> r = get_rows(std::vector<Row>{row});
        return std::move(r[0]);
    }

    std::vector<Row> get_column(Column column) const override {
        uint64_t n = 0;
        check(ggq_get_column_count(handle(), column, &n));   // host-side lookup
        std::vector<Row> out(n);
        check(ggq_get_column_into(handle(), column, out.data(), n, &n));   // straight into the result: one copy
        return out;
    }

    // ---- batched extensions ----------------------------------------------------------------
    std::vector<std::vector<Column>> get_rows(const std::vector<Row> &rows) const {
        ggq_rows_result *res = nullptr;
        check(ggq_get_rows(handle(), rows.data(), rows.size(), &res));
        const uint64_t *off = ggq_rows_offsets(res), *cols = ggq_rows_columns(res);
        std::vector<std::vector<Column>> out(rows.size());
        for (size_t t = 0; t < rows.size(); ++t) out[t].assign(cols + off[t], cols + off[t + 1]);
        ggq_rows_result_free(res);
        return out;
    }

    // CSR form of get_rows: offsets has rows.size() + 1 entries
    void get_rows_csr(const std::vector<Row> &rows, std::vector<uint64_t> *offsets, std::vector<Column> *cols) const {
        ggq_rows_result *res = nullptr;
        check(ggq_get_rows(handle(), rows.data(), rows.size(), &res));
        const uint64_t *off = ggq_rows_offsets(res);
        offsets->assign(off, off + rows.size() + 1);
        cols->assign(ggq_rows_columns(res), ggq_rows_columns(res) + off[rows.size()]);
        ggq_rows_result_free(res);
    }

    std::vector<Column> slice_rows(const std::vector<Row> &rows) const {
        uint64_t *buf = nullptr, n = 0;
        check(ggq_slice_rows(handle(), rows.data(), rows.size(), &buf, &n));
        std::vector<Column> out(buf, buf + n);
        ggq_buffer_free(buf);
        return out;
    }

    // Histogram of the set columns over a multiset of rows: the counter StaticBinRelAnnotator::count_labels fills
    // with one get_row per index (annotate_static.cpp:149-162), computed on the device without materialising rows.
    std::vector<uint64_t> count_columns(const std::vector<Row> &rows) const {
        std::vector<uint64_t> counts(num_columns());
        uint64_t dummy = 0;
        check(ggq_count_labels(handle(), rows.data(), rows.size(), counts.empty() ? &dummy : counts.data()));
        return counts;
    }

    // Columns kept by StaticBinRelAnnotator::get_labels(indices, presence_ratio) (annotate_static.cpp:71-94), ascending;
    // the caller decodes them with its LabelEncoder.  Histogram + threshold on the device.
    std::vector<Column> get_columns_with_ratio(const std::vector<Row> &rows, double presence_ratio) const {
        std::vector<Column> cols(std::max<uint64_t>(num_columns(), 1));
        uint64_t k = 0;
        check(ggq_get_labels(handle(), rows.data(), rows.size(), presence_ratio, cols.data(), nullptr, &k));
        cols.resize(k);
        return cols;
    }

    // (column, count) pairs of MultiLabelEncoded::get_top_labels(indices, num_top) (common/annotate.cpp:57-83):
    // decreasing count; equal counts in ascending column order (unspecified in the reference).
    std::vector<std::pair<Column, uint64_t>> get_top_columns(const std::vector<Row> &rows, size_t num_top) const {
        const uint64_t cap = std::max<uint64_t>(std::min<uint64_t>(num_columns(), num_top), 1);
        std::vector<uint64_t> cols(cap), counts(cap);
        uint64_t k = 0;
        check(ggq_get_top_labels(handle(), rows.data(), rows.size(), num_top, cols.data(), counts.data(), &k));
        std::vector<std::pair<Column, uint64_t>> out(k);
        for (uint64_t i = 0; i < k; ++i) out[i] = {cols[i], counts[i]};
        return out;
    }

    std::vector<bool> has_column(const std::vector<Row> &rows, Column column) const {
        std::vector<uint8_t> tmp(rows.size());
        check(ggq_has_column(handle(), rows.data(), rows.size(), column, tmp.data()));
        return std::vector<bool>(tmp.begin(), tmp.end());
    }

    // serialize() reproduces the stream the matrix was loaded from byte for byte (the device image is a different,
    // query-oriented layout).  By default load() keeps a host copy of that stream; a caller that loads from a file
    // can avoid the copy (a Kingsford-scale .brwt is ~10 GB): load_file(path), or keep_stream(false) +
    // set_stream_path(path), re-read the file on demand.  With neither a copy nor a path serialize() THROWS — it
    // never writes an empty or truncated matrix.
    void serialize(std::ostream &out) const override {
        if (!out.good()) throw std::ofstream::failure("Error when dumping the matrix");
        if (have_stream_) { out.write(stream_.data(), std::streamsize(stream_.size())); return; }
        if (stream_path_.empty()) {
            serialize_from_device(out);   // BRWT: re-encoded from the device image; others throw
            return;
        }
        std::ifstream in(stream_path_, std::ios::binary);
        if (!in.good()) throw std::ofstream::failure("ggq: cannot read the serialized matrix at " + stream_path_);
        in.seekg(std::streamoff(stream_offset_));
        std::vector<char> buf(size_t(1) << 20);
        uint64_t left = stream_bytes_;
        while (left) {
            in.read(buf.data(), std::streamsize(std::min<uint64_t>(left, buf.size())));
            const std::streamsize got = in.gcount();
            if (got <= 0) throw std::ofstream::failure("ggq: the serialized matrix at " + stream_path_ + " is shorter than when it was loaded");
            out.write(buf.data(), got);
            left -= uint64_t(got);
        }
    }

    // No host copy and no file: the engine writes the reference's stream format from the device image
    // (ggq_brwt_serialize; BRWT matrices that still have their host tree — loaded from a stream or built on the device).
    virtual void serialize_from_device(std::ostream &) const {
        throw std::ofstream::failure("ggq: serialize(): the matrix has neither a host copy of its stream nor a stream path");
    }

    // Do not keep a host copy of the stream read by the next load() (serialize() then needs set_stream_path()).
    void keep_stream(bool keep) { keep_stream_ = keep; }
    // Where serialize() re-reads the matrix from: `bytes` bytes at `offset` of `path` (e.g. behind the label encoder
    // of a .brwt.annodbg, annotate_static.cpp:96-108).
    void set_stream_path(const std::string &path, uint64_t offset, uint64_t bytes) {
        stream_path_ = path; stream_offset_ = offset; stream_bytes_ = bytes;
    }
    // load() from a file holding only the matrix stream, without the host copy.
    bool load_file(const std::string &path) {
        std::ifstream in(path, std::ios::binary);
        const bool keep = keep_stream_;
        keep_stream_ = false;
        const bool ok = in.good() && load(in);
        keep_stream_ = keep;
        if (ok) set_stream_path(path, 0, loaded_bytes_);
        return ok;
    }
    uint64_t loaded_stream_bytes() const { return loaded_bytes_; }

    int device() const { return device_; }
    uint64_t device_bytes() const { return h_ ? ggq_device_bytes(h_.get()) : 0; }
    const ggq_matrix *raw() const { return h_.get(); }

  protected:
    typedef int (*loader_fn)(ggq_read_fn, void *, int, ggq_matrix **);

    struct Tap {
        std::istream *in;
        std::string *copy;   // null: count only
        uint64_t bytes;
    };
    static size_t read_cb(void *ctx, void *dst, size_t n) {
        Tap *t = static_cast<Tap *>(ctx);
        t->in->read(static_cast<char *>(dst), std::streamsize(n));
        size_t got = size_t(t->in->gcount());
        if (t->copy) t->copy->append(static_cast<const char *>(dst), got);
        t->bytes += got;
        return got;
    }

    bool load_with(loader_fn fn, std::istream &in) {
        if (!in.good()) return false;
        try {
            std::string bytes;
            Tap tap{&in, keep_stream_ ? &bytes : nullptr, 0};
            ggq_matrix *m = nullptr;
            if (fn(&read_cb, &tap, device_, &m) != GGQ_OK) return false;
            h_.reset(m, [](ggq_matrix *p) { ggq_free(p); });
            stream_.swap(bytes);
            have_stream_ = keep_stream_;
            loaded_bytes_ = tap.bytes;
            stream_path_.clear();
            return true;
        } catch (...) {
            return false;
        }
    }

    const ggq_matrix *handle() const {
        if (!h_) throw std::runtime_error("ggq: matrix is not loaded");
        return h_.get();
    }
    static void check(int rc) {
        if (rc != GGQ_OK) throw std::runtime_error(std::string("ggq: ") + ggq_last_error());
    }

    int device_;
    std::shared_ptr<ggq_matrix> h_;   // copies share the immutable device image
    std::string stream_;              // host copy of the loaded stream (have_stream_)
    bool keep_stream_ = true, have_stream_ = false;
    std::string stream_path_;         // else: where serialize() re-reads it from
    uint64_t stream_offset_ = 0, stream_bytes_ = 0, loaded_bytes_ = 0;
};

class GpuBRWT final : public GpuMatrixBase {
  public:
    using GpuMatrixBase::GpuMatrixBase;
    bool load(std::istream &in) override { return load_with(&ggq_brwt_load, in); }
    // BRWT stats (BRWT.hpp:44-49)
    uint64_t num_nodes() const { return h_ ? ggq_num_nodes(h_.get()) : 0; }
    uint64_t total_column_size() const { return h_ ? ggq_total_column_size(h_.get()) : 0; }

    // Capacity layout (ggq_set_default_option "wide_slots"): matrices loaded / built AFTER the call use 32-byte slots of
    // 224 positions — 14 % instead of 33 % rank overhead, for matrices that do not fit a device otherwise; slower
    // descent (one sector per child), identical answers.  Process-wide.
    static void use_capacity_layout(bool on) {
        if (ggq_set_default_option("wide_slots", on ? 1 : 0) != GGQ_OK) throw std::runtime_error(ggq_last_error());
    }
    uint32_t slot_bits() const { return h_ ? ggq_slot_bits(h_.get()) : 0; }

    // Flattened device image (ggq_image_save / ggq_image_load): written once next to the .brwt.annodbg, it restores
    // the GPU-resident layout without parsing the stream again.  `stream_path` (optional) names the serialized
    // matrix the image was made from; it is read back only if serialize() is called.
    void save_image(const std::string &path) const { check(ggq_image_save(handle(), path.c_str())); }
    void serialize_from_device(std::ostream &out) const override {
        struct W {
            static size_t put(void *ctx, const void *src, size_t n) {
                std::ostream *o = static_cast<std::ostream *>(ctx);
                o->write(static_cast<const char *>(src), std::streamsize(n));
                return o->good() ? n : 0;
            }
        };
        if (ggq_brwt_serialize(handle(), &W::put, &out, nullptr) != GGQ_OK)
            throw std::ofstream::failure(std::string("ggq: serialize(): ") + ggq_last_error());
    }
    // Builds the matrix on the device from explicit columns (ggq_brwt_build_columns: the reference's bottom-up builder
    // with the basic or the greedy partitioner + relax); serialize() then writes what the reference builder would.
    bool build(const std::vector<std::vector<uint64_t>> &column_words, uint64_t n_rows, bool greedy, uint32_t arity = 2,
               uint64_t relax_arity = 0) {
        const uint64_t W = (n_rows + 63) / 64;
        std::vector<uint64_t> flat(column_words.size() * W);
        for (size_t j = 0; j < column_words.size(); ++j) {
            if (column_words[j].size() < W) return false;
            std::copy(column_words[j].begin(), column_words[j].begin() + W, flat.begin() + j * W);
        }
        ggq_matrix *m = nullptr;
        if (ggq_brwt_build_columns(flat.data(), n_rows, column_words.size(), greedy ? 1 : 0, arity, relax_arity, device_, &m) != GGQ_OK)
            return false;
        h_.reset(m, [](ggq_matrix *p) { ggq_free(p); });
        stream_.clear();
        have_stream_ = false;
        stream_path_.clear();
        return true;
    }
    // Image without a stream_path: serialize() throws (the image does not carry the host tree).
    bool load_image(const std::string &path, const std::string &stream_path = std::string()) {
        ggq_matrix *m = nullptr;
        if (ggq_image_load(path.c_str(), device_, &m) != GGQ_OK) return false;
        h_.reset(m, [](ggq_matrix *p) { ggq_free(p); });
        stream_.clear();
        have_stream_ = false;
        stream_path_.clear();
        if (!stream_path.empty()) {
            std::ifstream in(stream_path, std::ios::binary | std::ios::ate);
            if (!in.good()) return false;
            set_stream_path(stream_path, 0, uint64_t(in.tellg()));
        }
        return true;
    }
};

class GpuColMajor final : public GpuMatrixBase {
  public:
    using GpuMatrixBase::GpuMatrixBase;
    bool load(std::istream &in) override { return load_with(&ggq_colmajor_load, in); }
};

}  // namespace ggq

#endif  // GGQ_GPU_MATRIX_HPP_
