/* ggq — C ABI of the B200-native annotation-matrix query engine (libggq.so).
 *
 * This is the drop-in boundary for the hot path of ratschlab/genome_graph_annotation:
 * every entry point replaces the reference call cited beside it.  Plain pointers and sizes
 * only; no C++ types, no exceptions cross this boundary.  All functions return GGQ_OK (0) or a
 * negative ggq_status; ggq_last_error() gives the message for the calling thread.
 *
 * Host C++ adapters with the reference's own class interface (BinaryMatrix) are in
 * include/ggq/gpu_matrix.hpp; INTEGRATION.md shows how a maintainer wires them in.
 *
 * The library needs a CUDA device (sm_100a).  There is NO CPU fallback: without a usable
 * device every compute entry point fails with GGQ_ERR_CUDA.
 */
#ifndef GGQ_H_
#define GGQ_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum ggq_status {
    GGQ_OK = 0,
    GGQ_ERR_FORMAT = -1,   /* stream is not a valid serialized matrix (reference load() == false) */
    GGQ_ERR_CUDA = -2,     /* CUDA runtime failure, or no device */
    GGQ_ERR_OOM = -3,      /* host or device allocation failed */
    GGQ_ERR_ARG = -4,      /* invalid argument (null pointer, column out of range, ...) */
    GGQ_ERR_LIMIT = -5,    /* matrix exceeds an engine limit (see DESIGN.md "Limits") */
    GGQ_ERR_INTERNAL = -6,
    GGQ_ERR_NCCL = -7      /* NCCL failure in a column-sharded query */
} ggq_status;

typedef struct ggq_matrix ggq_matrix;            /* GPU-resident BinaryMatrix (BRWT or column-major) */
typedef struct ggq_bv ggq_bv;                    /* GPU-resident bit vector with rank/select */
typedef struct ggq_workspace ggq_workspace;      /* per-caller stream + scratch (device-resident API) */
typedef struct ggq_comm ggq_comm;                /* NCCL communicator of the shards of one matrix */
typedef struct ggq_rows_result ggq_rows_result;  /* host-resident CSR result of ggq_get_rows */

/* Pull reader used by the stream loaders: must copy exactly n bytes to dst and return n, or return
 * fewer on EOF/error.  Lets BinaryMatrix::load(std::istream&) consume exactly the matrix' bytes. */
typedef size_t (*ggq_read_fn)(void *ctx, void *dst, size_t n);
/* Output callback of ggq_brwt_serialize: takes n bytes, returns the number it took (fewer than n = failure). */
typedef size_t (*ggq_write_fn)(void *ctx, const void *src, size_t n);

const char *ggq_last_error(void);
int ggq_device_count(int *count);

/* ------------------------------------------------------------------ load / upload ------------ */

/* BRWT::load (annotation/hierarchical_annotation/BRWT.cpp:87-111): parses the reference's
 * serialized BRWT / Multi-BRWT stream (pre-order nodes: RangePartition, rrr_vector<63>, children),
 * flattens it into level-order sibling-interleaved arrays and uploads it once to `device`. */
int ggq_brwt_load(ggq_read_fn read, void *ctx, int device, ggq_matrix **out);
int ggq_brwt_load_buffer(const uint8_t *buf, size_t len, size_t *consumed, int device, ggq_matrix **out);
/* BRWT::load from a file holding the matrix stream (buffered stdio; a missing file: GGQ_ERR_FORMAT, as load() == false).
 * Host memory during the load: the compressed vectors (about the file size) + at most 1 GiB of decoded bits. */
int ggq_brwt_load_file(const char *path, int device, ggq_matrix **out);

/* ColMajorCompressed::load (annotation/column_compressed/column_major.cpp:58-79): BE u64 column
 * count + bit_vector_sd streams. */
int ggq_colmajor_load(ggq_read_fn read, void *ctx, int device, ggq_matrix **out);
int ggq_colmajor_load_buffer(const uint8_t *buf, size_t len, size_t *consumed, int device, ggq_matrix **out);

/* Benchmark harness: builds ON THE DEVICE the BRWT that
 * BRWTBottomUpBuilder::build(columns, get_basic_partitioner(arity)) (BRWT_builders.cpp:19-31,
 * 122-163) would produce for the implicit matrix  M[i][j] = mix64(seed*0x9E3779B97F4A7C15 + i*n_cols + j) < threshold. */
int ggq_brwt_synthesize(uint64_t n_rows, uint64_t n_cols, uint64_t threshold, uint64_t seed,
                        uint32_t arity, int device, ggq_matrix **out);

/* Flattened device-image file.  ggq_image_save writes the GPU-resident layout of a BRWT (header,
 * geometry tables with a checksum, slot arena, select samples) to `path`; ggq_image_load uploads it as it is —
 * no RRR decode, no flattening, no rank-header construction (what BRWT::load + upload cost at every start,
 * BRWT.cpp:87-111).  The loaded matrix answers every query identically; BinaryMatrix::serialize still needs the
 * original stream (the adapters keep it).  A truncated, foreign-endian or other-version file: GGQ_ERR_FORMAT. */
int ggq_image_save(const ggq_matrix *m, const char *path);
int ggq_image_load(const char *path, int device, ggq_matrix **out);

void ggq_free(ggq_matrix *m);

/* Host-only: parse + flatten a serialized BRWT exactly as ggq_brwt_load does, without touching the
 * device (works on a box with no GPU; used by the CPU tests and by tools that size a deployment).
 * stats[0..7] = num_rows, num_columns, num_nodes, num_relations, total_column_size (BRWT.hpp:44-49),
 * sibling groups, device image bytes (in the slot layout ggq_set_default_option currently selects), max depth. */
int ggq_brwt_inspect_buffer(const uint8_t *buf, size_t len, size_t *consumed, uint64_t stats[8]);
/* Host-only: DFS (= get_row) order of the columns and the root-to-leaf depth of every column of a
 * serialized BRWT.  dfs_cols / depth: num_columns entries each, caller-allocated (may be NULL). */
int ggq_brwt_inspect_columns(const uint8_t *buf, size_t len, uint32_t *dfs_cols, uint32_t *depth);

/* BRWT::serialize (annotation/hierarchical_annotation/BRWT.cpp:113-128) FROM THE DEVICE IMAGE: every node's vector is
 * read back from the arena and re-encoded as the reference's rrr_vector<63>, byte for byte what the reference writes
 * for the same tree — so a matrix built on the device (ggq_brwt_build_columns) becomes a .brwt the reference loads, and
 * a loaded matrix needs no host copy of its stream to be written back (matrices restored from an image file included:
 * the image carries the host tree).  Not available for column-major matrices and column shards (GGQ_ERR_INTERNAL). */
int ggq_brwt_serialize(const ggq_matrix *m, ggq_write_fn write, void *ctx, uint64_t *n_bytes);
int ggq_brwt_serialize_buffer(const ggq_matrix *m, uint8_t **out, size_t *len);   /* *out: ggq_buffer_free */
/* The same through buffered stdio, for streams of many gigabytes: no copy of the stream in host memory.  A file that
 * cannot be created: GGQ_ERR_ARG; a failed write: GGQ_ERR_INTERNAL. */
int ggq_brwt_serialize_file(const ggq_matrix *m, const char *path, uint64_t *n_bytes);

/* Host only: the bytes sdsl's rrr_vector<63>(bits).serialize() writes for the given plain bits (LSB-first words) —
 * the building block of ggq_brwt_serialize.  *out is malloc'ed (ggq_buffer_free). */
int ggq_rrr63_encode_buffer(const uint64_t *words, uint64_t n_bits, uint8_t **out, size_t *len);
/* Host only: the inverse, by the decoder BRWT::load uses here (sdsl rrr_vector.hpp:101-109 load, rrr_helper.hpp
 * decode): *n_bits = length of the vector; its bits go to `words` (LSB first, capacity n_words; too small: GGQ_ERR_ARG
 * with *n_bits set).  A damaged vector: GGQ_ERR_FORMAT. */
int ggq_rrr63_decode_buffer(const uint8_t *buf, size_t len, size_t *consumed, uint64_t *words, uint64_t n_words,
                            uint64_t *n_bits);

/* The reference's BRWT construction on the device: BRWTBottomUpBuilder::build(columns, partitioner)
 * (annotation/hierarchical_annotation/BRWT_builders.cpp:122-163) with partitioner 0 = get_basic_partitioner(arity)
 * (:19-31) or 1 = parallel_binary_grouping_greedy (partitionings.cpp:146-196: pairs ranked by popcount(a & b) over
 * up to 10^6 sampled rows), followed by BRWTOptimizer::relax(relax_arity) (BRWT_builders.cpp:166-380) when
 * relax_arity > 1.  col_words: n_cols columns of ceil(n_rows / 64) LSB-first u64 words each.  Produces the SAME tree
 * as the reference builder (nodes, child order, column arrangement): the Gram matrix, the ORs and the index
 * restriction run on the GPU, the ranking and the space model decide as the reference does. */
int ggq_brwt_build_columns(const uint64_t *col_words, uint64_t n_rows, uint64_t n_cols, int partitioner, uint32_t arity,
                           uint64_t relax_arity, int device, ggq_matrix **out);
/* Benchmark harness: the same builder (greedy partitioner + relax(relax_arity)) for the IMPLICIT matrix of
 * ggq_brwt_synthesize, at sizes where the columns cannot be materialised — the pairing looks at the sampled rows only
 * (generated on the device), relax at popcounts counted over all rows; same tree as ggq_brwt_build_columns on the
 * materialised columns. */
int ggq_brwt_synthesize_greedy(uint64_t n_rows, uint64_t n_cols, uint64_t threshold, uint64_t seed, uint64_t relax_arity,
                               int device, ggq_matrix **out);
/* dfs_cols[k] = the column whose leaf is visited k-th by BRWT::get_row (DFS over child_nodes_, BRWT.cpp:45-50): the
 * order in which every row lists its columns.  num_columns entries. */
int ggq_tree_dfs_columns(const ggq_matrix *m, uint32_t *dfs_cols);
/* Tree of a BRWT matrix, nodes in level (BFS) order, children in stored order: parent index (-1: root), number of
 * children / columns, length and popcount of the node's vector, first column of its arrangement.  Arrays may be NULL;
 * *n_nodes is always set.  Not available for column-major matrices and images of column shards (GGQ_ERR_ARG). */
int ggq_tree_shape(const ggq_matrix *m, uint64_t *n_nodes, int64_t *parent, uint32_t *n_children, uint32_t *n_columns,
                   uint64_t *size, uint64_t *ones, uint32_t *first_column);

/* BinaryMatrix::num_rows / num_columns / num_relations (common/binary_matrix.hpp:16-28) */
uint64_t ggq_num_rows(const ggq_matrix *m);
uint64_t ggq_num_columns(const ggq_matrix *m);
uint64_t ggq_num_relations(const ggq_matrix *m);
/* BRWT stats (BRWT.hpp:44-49); 0 for column-major */
uint64_t ggq_num_nodes(const ggq_matrix *m);
uint64_t ggq_total_column_size(const ggq_matrix *m);
uint64_t ggq_device_bytes(const ggq_matrix *m);
/* Positions per slot of a BRWT image: 96 (default layout) or 224 (capacity layout, below); 0 for column-major. */
uint32_t ggq_slot_bits(const ggq_matrix *m);
/* Process-wide defaults for the matrices created AFTER the call (load, build, synthesize; an image file keeps the
 * layout it was written with).  "wide_slots": 0 = 16-byte slots {rank, 96 bits} (default; fastest: one 32-byte sector
 * answers two children), 1 = the capacity layout: 32-byte slots {rank, 224 bits} — 14 % instead of 33 % rank-header
 * overhead (a matrix that does not fit in the default layout may fit in this one), one sector per child;
 * vectors are limited to 2^32 positions.  Results are identical in both layouts.  Unknown name: GGQ_ERR_ARG. */
int ggq_set_default_option(const char *name, uint64_t value);
/* Leaf codes of a BRWT image: what[0] = coded nodes, what[1] = code bytes on the device,
 * what[2] = sampled escape rate x 1e6 (visit-weighted), what[3] = 1 if the descent currently uses them. */
int ggq_leaf_code_info(const ggq_matrix *m, uint64_t what[4]);
/* Algorithmic work of the last ggq_get_rows* call on this workspace/matrix: number of sibling-group
 * probes (one 16/32/64-byte block read each).  For roofline accounting. */
uint64_t ggq_last_probe_count(const ggq_matrix *m);

/* Tuning knobs of the row-query path (defaults are chosen per matrix at load time; DESIGN.md "K1").  Call
 * before creating workspaces on the matrix.  Names: "rows_per_warp" (1..16), "sort_bits" (0 = process queries in
 * the caller's order; k = order each batch by the top k bits of the row id), "sort_min_rows" (smaller batches are
 * never reordered).  Results never depend on these. */
int ggq_set_option(ggq_matrix *m, const char *name, uint64_t value);

/* ------------------------------------------------------------------ row queries -------------- */

/* get_rows(R)[t] == BRWT::get_row(R[t]) (BRWT.cpp:26-53) / ColMajorCompressed::get_row
 * (column_major.cpp:34-47): for every row id the set columns in the reference's order (DFS over
 * children for BRWT, ascending for column-major).  Host buffers in, host CSR out.
 * Row ids >= num_rows (undefined behaviour in the reference, assert only) yield empty rows. */
int ggq_get_rows(const ggq_matrix *m, const uint64_t *rows, uint64_t n, ggq_rows_result **out);
const uint64_t *ggq_rows_offsets(const ggq_rows_result *r);  /* n + 1 entries */
const uint64_t *ggq_rows_columns(const ggq_rows_result *r);  /* offsets[n] entries */
uint64_t ggq_rows_count(const ggq_rows_result *r);
void ggq_rows_result_free(ggq_rows_result *r);

/* Same query, caller-provided host output (cols_capacity entries).  *n_cols receives the number of
 * column ids; if it exceeds cols_capacity GGQ_ERR_ARG is returned (offsets are complete, the contents of
 * cols are unspecified).  Batches above 2^20 rows are pipelined in chunks over three streams: a chunk is enqueued
 * before the host waits for the size of the previous one, and the device->host copy of a chunk overlaps the
 * descent of the next ones; page-locked (pinned) caller buffers are needed for that overlap, pageable ones work
 * but serialise. */
int ggq_get_rows_into(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t *offsets,
                      uint64_t *cols, uint64_t cols_capacity, uint64_t *n_cols);

/* Compact variant: column ids as u32 (the engine holds < 2^31 columns, so nothing is lost) — halves the
 * device->host traffic, which is what bounds ggq_get_rows_into on PCIe (DESIGN.md "Measurement"). */
int ggq_get_rows_into_u32(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t *offsets,
                          uint32_t *cols, uint64_t cols_capacity, uint64_t *n_cols);

/* Compact form of the same query for callers bound by the host link (8 GPUs returning every row set to one host:
 * DESIGN.md "Measurement").  Row ids of row_bytes = 4 (matrices of at most 2^32 rows) or 8; per-row COUNTS instead
 * of offsets, count_bytes = 1, 2 or 4 each (row t's ids start at the sum of the counts before it); column ids of
 * col_bytes = 2 (matrices of at most 65536 columns), 4 or 8, rows in the caller's order, ids in get_row order.
 * Same results as ggq_get_rows_into, 2.5x fewer bytes on the C4 matrix (22 instead of 56 per row).
 * A row with more ids than count_bytes can hold: GGQ_ERR_LIMIT (counts saturated; call again with wider counts). */
int ggq_get_rows_compact(const ggq_matrix *m, const void *rows, int row_bytes, uint64_t n, void *counts, int count_bytes,
                         void *cols, int col_bytes, uint64_t cols_capacity, uint64_t *n_cols);

/* counts[j] = number of t with column j set in row rows[t] (duplicates counted) — the histogram
 * StaticBinRelAnnotator::count_labels builds with one get_row per index (annotation/annotate_static.cpp:149-162),
 * i.e. what get_labels(indices, ratio) / get_top_labels threshold (annotate_static.cpp:71-94, annotate.cpp:57-83).
 * The row sets never leave the device: descent + shared-memory histogram, num_columns counters come back.
 * counts: num_columns entries.  Column-sharded matrices count their owned columns (others stay 0: sum over ranks). */
int ggq_count_labels(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t *counts);
int ggq_count_labels_device(const ggq_matrix *m, ggq_workspace *ws, const uint64_t *d_rows, uint64_t n,
                            const uint64_t **d_counts);

/* StaticBinRelAnnotator::get_labels(indices, presence_ratio) (annotation/annotate_static.cpp:71-94): the columns with
 * counts[j] != 0 && counts[j] >= (ratio == 0 ? 1 : ceil(n * ratio)), ascending, with their counts.  Histogram and
 * threshold run on the device; only the kept pairs come back.  cols / counts: num_columns entries (either may be
 * NULL); *n_out receives the number of kept columns.  presence_ratio outside [0, 1] (assert in the reference):
 * GGQ_ERR_ARG. */
int ggq_get_labels(const ggq_matrix *m, const uint64_t *rows, uint64_t n, double presence_ratio, uint64_t *cols,
                   uint64_t *counts, uint64_t *n_out);
/* MultiLabelEncoded::get_top_labels(indices, num_top) (common/annotate.cpp:57-83): the non-zero (column, count) pairs
 * in decreasing order of count, truncated to num_top.  The reference's std::sort leaves the order of equal counts
 * unspecified; here ties are in ascending column order (stable device radix sort).  cols / counts: min(num_top,
 * num_columns) entries. */
int ggq_get_top_labels(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t num_top, uint64_t *cols,
                       uint64_t *counts, uint64_t *n_out);

/* Cross-query batching: `annograph classify` runs one execute_query task per sequence
 * (main.cpp:462-504), each calling get_labels(indices, ratio) on 10^2..10^3 indices — too few to fill a GPU.  Here the
 * index lists of n_queries sequences are concatenated (query q owns rows[query_offsets[q] .. query_offsets[q+1]),
 * query_offsets[0] == 0) and answered by ONE descent launch plus one histogram CTA per query.  For every query the
 * result is exactly ggq_get_labels(rows of q, presence_ratio): label_offsets has n_queries + 1 entries, query q's
 * columns (ascending) and counts are cols/counts[label_offsets[q] .. label_offsets[q+1]).  capacity: entries of
 * cols / counts (counts may be NULL); *n_out = total pairs; GGQ_ERR_ARG if it exceeds capacity (label_offsets is
 * complete).  Matrices with more columns than fit a shared-memory histogram (~57k): GGQ_ERR_LIMIT. */
int ggq_get_labels_batch(const ggq_matrix *m, const uint64_t *rows, const uint64_t *query_offsets, uint64_t n_queries,
                         double presence_ratio, uint64_t *label_offsets, uint64_t *cols, uint64_t *counts,
                         uint64_t capacity, uint64_t *n_out);
/* Same with one presence ratio PER QUERY — what AnnotatedDBG::get_labels passes: it rescales the caller's ratio by
 * the k-mers of the sequence that are missing from the graph (common/annotated_dbg.cpp:104-113).  Batches of up to
 * 2^14 indices in all take a direct form: one host->device copy, descent, one CTA per query that builds the
 * histogram from the descent's chunks and writes the kept pairs into pinned host memory, one wait. */
int ggq_get_labels_batch_ratios(const ggq_matrix *m, const uint64_t *rows, const uint64_t *query_offsets, uint64_t n_queries,
                                const double *presence_ratios, uint64_t *label_offsets, uint64_t *cols, uint64_t *counts,
                                uint64_t capacity, uint64_t *n_out);

/* slice_rows: concatenation of get_row(R[t]) each followed by UINT64_MAX (SURVEY.md "Three facts" 1). */
int ggq_slice_rows(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t **out, uint64_t *out_len);
void ggq_buffer_free(void *p);

/* Device-resident variant: d_rows is device memory on the matrix' device; the CSR stays on the
 * device in buffers owned by the workspace (valid until the next call on that workspace).
 * col_bytes = 4 or 8 selects u32 / u64 column ids.  Runs on the workspace's stream; synchronizes
 * once to learn the result size. */
int ggq_workspace_create(const ggq_matrix *m, ggq_workspace **out);
void ggq_workspace_free(ggq_workspace *ws);
void *ggq_workspace_stream(ggq_workspace *ws);  /* cudaStream_t */
/* Measurements of the last query on this workspace (for roofline accounting, DESIGN.md "Measurement"):
 * device time of the dominant kernel between two CUDA events on the workspace stream; sibling-group
 * probes (one 16/32/64-byte block each); (node,row) bit tests (V_t of SURVEY.md 8d); kernel launches. */
enum { GGQ_STAT_KERNEL_MS = 0, GGQ_STAT_PROBES = 1, GGQ_STAT_TOUCHES = 2, GGQ_STAT_LAUNCHES = 3,
       GGQ_STAT_EXCHANGE_MS = 4 /* NCCL exchange of the last sharded query */, GGQ_STAT_EXCHANGE_BYTES = 5,
       GGQ_STAT_CODE_PROBES = 6 /* 1-byte leaf-code probes of the last query */, GGQ_STAT_CODE_ESCAPES = 7 };
double ggq_workspace_stat(const ggq_workspace *ws, int what);
int ggq_get_rows_device(const ggq_matrix *m, ggq_workspace *ws, const uint64_t *d_rows, uint64_t n,
                        int col_bytes, const uint64_t **d_offsets, const void **d_cols, uint64_t *n_cols);

/* ------------------------------------------------------------------ column sharding ---------- */
/* SURVEY.md 8(e) "root-subtree column sharding" (config C5: the image does not fit one GPU).  The tree is cut at
 * the smallest depth whose frontier has >= 4 elements per shard; the frontier is dealt to shards in contiguous
 * DFS runs balanced by column count (every rank derives the same plan from the tree alone).  Levels above the
 * cut are replicated, a deeper subtree lives only on its owner.  A shard answers get_rows with the owned columns
 * only; concatenating the shards' rows in shard order IS BRWT::get_row's order (BRWT.cpp:45-50).
 * get / has_column / get_column work on the owning shard and fail with GGQ_ERR_ARG elsewhere;
 * num_relations counts owned columns only. */
int ggq_brwt_load_shard(ggq_read_fn read, void *ctx, int device, uint32_t shard, uint32_t n_shards, ggq_matrix **out);
int ggq_brwt_load_shard_buffer(const uint8_t *buf, size_t len, size_t *consumed, int device, uint32_t shard,
                               uint32_t n_shards, ggq_matrix **out);
int ggq_brwt_synthesize_shard(uint64_t n_rows, uint64_t n_cols, uint64_t threshold, uint64_t seed, uint32_t arity,
                              int device, uint32_t shard, uint32_t n_shards, ggq_matrix **out);
/* Host-only: the plan for a serialized BRWT.  col_owner: num_columns entries (may be NULL). */
int ggq_brwt_shard_plan_buffer(const uint8_t *buf, size_t len, uint32_t n_shards, uint32_t *col_owner, uint32_t *cut_depth);
/* col_owned: num_columns bytes (may be NULL) */
int ggq_shard_info(const ggq_matrix *m, uint32_t *shard, uint32_t *n_shards, uint8_t *col_owned);

/* Merge step alone: per query, concatenate the shards' partial rows in shard order.  d_offsets[s] (n + 1 u64) and
 * d_cols[s] (u32 ids) are device pointers readable from the workspace's device — local buffers filled by the
 * caller's own collective, or peer memory.  Result in buffers owned by the workspace. */
int ggq_merge_shards_device(ggq_workspace *ws, uint64_t n, int n_shards, const uint64_t *const *d_offsets,
                            const uint32_t *const *d_cols, int col_bytes, const uint64_t **d_out_offsets,
                            const void **d_out_cols, uint64_t *n_cols);

/* NCCL plumbing: rank 0 calls ggq_comm_unique_id and ships the 128 bytes to the other ranks by any means; every
 * rank (one process per GPU) then calls ggq_comm_init — collective. */
int ggq_comm_unique_id(uint8_t id[128]);
int ggq_comm_init(const uint8_t id[128], int rank, int n_ranks, int device, ggq_comm **out);
void ggq_comm_free(ggq_comm *c);
/* Peer-memory exchange — collective, optional, one process per GPU on one NVLink/NVSwitch node.  Pins the partial-CSR
 * buffers of `ws` at the given capacities (rows per batch, partial column ids per shard and batch), exports them with
 * cudaIpc and maps every peer's.  ggq_get_rows_sharded_device on that workspace then skips the NCCL all-gather: its
 * merge kernels read the shards' segments directly from peer memory over NVLink (transfer and merge are one pass),
 * with two 8-byte all-gathers as barriers.  A batch beyond the capacities falls back to the NCCL exchange for good
 * (every rank takes the same decision).  ggq_comm_peer_active tells which path the next query takes. */
int ggq_comm_enable_peer(ggq_comm *c, ggq_workspace *ws, uint64_t max_rows, uint64_t max_ids);
int ggq_comm_peer_active(const ggq_comm *c);
/* Collective over the shards of one matrix: every rank passes the SAME batch (device memory) and ends with the
 * complete CSR: local descent, ncclAllGather of partial offsets + grouped ncclBroadcast of partial column ids over
 * NVLink, merge kernel.  Runs on the workspace stream. */
int ggq_get_rows_sharded_device(const ggq_matrix *m, ggq_workspace *ws, ggq_comm *comm, const uint64_t *d_rows,
                                uint64_t n, int col_bytes, const uint64_t **d_offsets, const void **d_cols,
                                uint64_t *n_cols);
/* Same collective, but rank k merges only ITS slice of the batch: queries [n*k/R, n*(k+1)/R) (returned in
 * *first_query / *n_queries; d_offsets has *n_queries + 1 entries starting at 0).  Every rank still descends the
 * whole batch over its own columns; what shrinks is the exchange — with the peer-memory merge a rank reads 1/R of the
 * other shards' partial rows over NVLink instead of all of them. */
int ggq_get_rows_sharded_slice_device(const ggq_matrix *m, ggq_workspace *ws, ggq_comm *comm, const uint64_t *d_rows,
                                      uint64_t n, int col_bytes, uint64_t *first_query, uint64_t *n_queries,
                                      const uint64_t **d_offsets, const void **d_cols, uint64_t *n_cols);

/* count_labels across GPUs — collective.  Replicas: every rank passes ITS slice of the index list; column shards:
 * every rank passes the whole list (a shard counts its own columns, the others stay 0).  Either way one
 * ncclAllReduce(sum) over num_columns counters leaves the histogram of the whole batch on every rank — 21 KB over
 * NVLink instead of the row sets. */
int ggq_count_labels_allreduce_device(const ggq_matrix *m, ggq_workspace *ws, ggq_comm *comm, const uint64_t *d_rows,
                                      uint64_t n, const uint64_t **d_counts);

/* ------------------------------------------------------------------ cells / columns ---------- */

/* out[t] = BinaryMatrix::get(rows[t], cols[t]) (BRWT.cpp:9-24, column_major.cpp:27-32). */
int ggq_get(const ggq_matrix *m, const uint64_t *rows, const uint64_t *cols, uint64_t n, uint8_t *out);
/* has_column(R, j)[t] == get(R[t], j) */
int ggq_has_column(const ggq_matrix *m, const uint64_t *rows, uint64_t n, uint64_t col, uint8_t *out);
/* BinaryMatrix::get_column (BRWT.cpp:55-85, column_major.cpp:49-56): ascending row ids.
 * *rows is malloc'ed by the engine; release with ggq_buffer_free. */
int ggq_get_column(const ggq_matrix *m, uint64_t col, uint64_t **rows, uint64_t *n_rows);
/* Same with a caller-provided buffer (page-locked memory makes the copy run at PCIe speed: on the C4 matrix the
 * device needs 1.5 ms for a 1.1e7-row column, a pageable malloc'ed result costs ~100x that in page faults).
 * ggq_get_column_count is a host-side lookup (popcount of the column's leaf vector); `capacity` in entries. */
int ggq_get_column_count(const ggq_matrix *m, uint64_t col, uint64_t *n_rows);
int ggq_get_column_into(const ggq_matrix *m, uint64_t col, uint64_t *rows, uint64_t capacity, uint64_t *n_rows);
/* ggq_get_column_into for callers that may not allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The result buffer the K3 kernels work in is a fixed device array inside the library (8 MiB, held by every GPU that loads
 * the library until exit) instead of a workspace buffer: the call performs no cudaMalloc / cudaFree and calls no allocating
 * library routine.  BRWT matrices only (both slot layouts).  A column with more than GGQ_NOALLOC_MAX_COLUMN_ROWS set rows:
 * GGQ_ERR_LIMIT, *n_rows = the count (use ggq_get_column_into); `capacity` too small: GGQ_ERR_ARG.  Runs on the default
 * stream and returns when `rows` (host memory) is filled.  All calls share the one array per device: calls of one process
 * are serialised inside the library, but work that overlaps on the GPU by other means (several processes' captured graphs
 * replayed together, a caller's own streams reading the array) races.  Same answer as ggq_get_column_into. */
#define GGQ_NOALLOC_MAX_COLUMN_ROWS (1u << 20)
int ggq_get_column_noalloc(const ggq_matrix *m, uint64_t col, uint64_t *rows, uint64_t capacity, uint64_t *n_rows);

/* ------------------------------------------------------------------ bit vectors -------------- */

enum { GGQ_BV_STAT = 0, GGQ_BV_SD = 1, GGQ_BV_RRR = 2 };

/* bit_vector_{stat,sd,rrr<63>}::load (common/bit_vector.cpp:360-385, 725-745, 906-925). */
int ggq_bv_load_buffer(int kind, const uint8_t *buf, size_t len, size_t *consumed, int device, ggq_bv **out);
/* Same device layouts from plain bits (LSB-first u64 words), as the reference's constructors from
 * sdsl::bit_vector do (common/bit_vector.cpp:253-262, 572-590, 806-812). */
int ggq_bv_build(int kind, const uint64_t *words, uint64_t n_bits, int device, ggq_bv **out);
/* Benchmark harness (config C3): bit i set iff mix64(seed*0x9E3779B97F4A7C15 + i) < threshold, generated on device. */
int ggq_bv_synthesize(int kind, uint64_t n_bits, uint64_t threshold, uint64_t seed, int device, ggq_bv **out);
void ggq_bv_free(ggq_bv *v);
uint64_t ggq_bv_size(const ggq_bv *v);
uint64_t ggq_bv_num_set_bits(const ggq_bv *v);
int ggq_bv_is_inverted(const ggq_bv *v);  /* bit_vector_sd::is_inverted (bit_vector.hpp:171) */
uint64_t ggq_bv_device_bytes(const ggq_bv *v);

/* rank1: ones in [0, id], id >= size clamps (bit_vector.hpp:16-17, bit_vector.cpp:285,692,860).
 * select1: position of the i-th one, i in [1, num_set_bits] (bit_vector.hpp:19-20).
 * get: operator[].  Host arrays; `device` variants take device pointers and a cudaStream_t. */
int ggq_bv_rank1(const ggq_bv *v, const uint64_t *q, uint64_t n, uint64_t *out);
int ggq_bv_select1(const ggq_bv *v, const uint64_t *q, uint64_t n, uint64_t *out);
int ggq_bv_get(const ggq_bv *v, const uint64_t *q, uint64_t n, uint8_t *out);
int ggq_bv_rank1_device(const ggq_bv *v, const uint64_t *d_q, uint64_t n, uint64_t *d_out, void *stream);
int ggq_bv_select1_device(const ggq_bv *v, const uint64_t *d_q, uint64_t n, uint64_t *d_out, void *stream);
int ggq_bv_get_device(const ggq_bv *v, const uint64_t *d_q, uint64_t n, uint8_t *d_out, void *stream);
/* The rest of the reference's bit_vector interface (not on the row-query path; composed from rank1 / operator[] on
 * the device, as the reference composes them on the host):
 *   rank0(id) = min(id + 1, size) - rank1(id)                         (common/bit_vector.cpp:32-34)
 *   select0(k): position of the k-th zero, k in [1, size - num_set_bits] (bit_vector_rrr::select0, bit_vector.hpp:199;
 *               answered for every kind, by bisection over rank1)
 *   get_int(id, width): bits [id, id + width) LSB first, width in [1, 64] (bit_vector.hpp:25); bits past the end read 0
 *   to_words: the plain bits, (size + 63) / 64 LSB-first words         (to_vector / copy, bit_vector.cpp:36-46)
 *   call_ones_into: positions of the set bits, ascending               (call_ones, bit_vector.cpp:442-451, 767-790, 958-983);
 *               *n_ones is always set; GGQ_ERR_ARG if capacity < num_set_bits */
/* AnnotatedDBG::graph_to_anno_index (common/annotated_dbg.cpp:142-156) for a batch: anno_idx[t] =
 * annotation_mask.rank1(graph_idx[t]) - 1 where the mask bit is set.  Where it is not (the reference throws "trying
 * to annotate k-mer that cannot be annotated") or the index is out of range, anno_idx[t] = UINT64_MAX — a value every
 * row query answers with an empty row — and *n_unmapped (may be NULL) counts them.  The device variant chains into
 * ggq_get_rows_device / ggq_count_labels_device on the same stream without a host round trip. */
int ggq_map_rank1(const ggq_bv *mask, const uint64_t *graph_idx, uint64_t n, uint64_t *anno_idx, uint64_t *n_unmapped);
int ggq_map_rank1_device(const ggq_bv *mask, const uint64_t *d_graph_idx, uint64_t n, uint64_t *d_anno_idx, void *stream);
int ggq_bv_rank0(const ggq_bv *v, const uint64_t *q, uint64_t n, uint64_t *out);
int ggq_bv_select0(const ggq_bv *v, const uint64_t *q, uint64_t n, uint64_t *out);
int ggq_bv_rank0_device(const ggq_bv *v, const uint64_t *d_q, uint64_t n, uint64_t *d_out, void *stream);
int ggq_bv_select0_device(const ggq_bv *v, const uint64_t *d_q, uint64_t n, uint64_t *d_out, void *stream);
int ggq_bv_get_int(const ggq_bv *v, const uint64_t *q, uint64_t n, uint32_t width, uint64_t *out);
int ggq_bv_to_words(const ggq_bv *v, uint64_t *words);
int ggq_bv_call_ones_into(const ggq_bv *v, uint64_t *positions, uint64_t capacity, uint64_t *n_ones);

#ifdef __cplusplus
}
#endif
#endif /* GGQ_H_ */
