"""B200-native query engine for genome-graph annotation matrices (BRWT / Multi-BRWT / column-major).

Python is only a thin ctypes binding over the C ABI of ``libggq.so`` (include/ggq.h) — used by the
tests and by bench.py.  The product is the CUDA library; the reference-facing host interface is the
C++ ``BinaryMatrix`` adapter in include/ggq/gpu_matrix.hpp.

There is no CPU fallback: importing works without a GPU (so that the symbol table can be checked),
but every compute call needs a CUDA device and raises :class:`GgqError` otherwise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libggq.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "ggq.h")

u64p = C.POINTER(C.c_uint64)
u8p = C.POINTER(C.c_uint8)
READ_FN = C.CFUNCTYPE(C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t)

GGQ_OK = 0
STATUS = {0: "GGQ_OK", -1: "GGQ_ERR_FORMAT", -2: "GGQ_ERR_CUDA", -3: "GGQ_ERR_OOM", -4: "GGQ_ERR_ARG",
          -5: "GGQ_ERR_LIMIT", -6: "GGQ_ERR_INTERNAL"}


class GgqError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("%s: %s" % (STATUS.get(code, code), msg))
        self.code = code


_lib = None


def lib():
    """Load libggq.so (built in-tree by ``make -C genome_graph_annotation_b200/csrc``).  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libggq.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "or `make -C genome_graph_annotation_b200/csrc` (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    sig = {
        "ggq_last_error": (C.c_char_p, []),
        "ggq_device_count": (C.c_int, [C.POINTER(C.c_int)]),
        "ggq_brwt_load": (C.c_int, [READ_FN, vp, C.c_int, C.POINTER(vp)]),
        "ggq_brwt_load_buffer": (C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t), C.c_int, C.POINTER(vp)]),
        "ggq_colmajor_load": (C.c_int, [READ_FN, vp, C.c_int, C.POINTER(vp)]),
        "ggq_colmajor_load_buffer": (C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t), C.c_int, C.POINTER(vp)]),
        "ggq_brwt_synthesize": (C.c_int, [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_int, C.POINTER(vp)]),
        "ggq_free": (None, [vp]),
        "ggq_brwt_load_shard": (C.c_int, [READ_FN, vp, C.c_int, C.c_uint32, C.c_uint32, C.POINTER(vp)]),
        "ggq_brwt_load_shard_buffer": (C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t), C.c_int, C.c_uint32, C.c_uint32, C.POINTER(vp)]),
        "ggq_brwt_synthesize_shard": (C.c_int, [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32, C.POINTER(vp)]),
        "ggq_brwt_shard_plan_buffer": (C.c_int, [C.c_char_p, C.c_size_t, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
        "ggq_shard_info": (C.c_int, [vp, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), u8p]),
        "ggq_merge_shards_device": (C.c_int, [vp, C.c_uint64, C.c_int, C.POINTER(vp), C.POINTER(vp), C.c_int, C.POINTER(vp), C.POINTER(vp), u64p]),
        "ggq_comm_unique_id": (C.c_int, [u8p]),
        "ggq_comm_init": (C.c_int, [u8p, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]),
        "ggq_comm_free": (None, [vp]),
        "ggq_count_labels_allreduce_device": (C.c_int, [vp, vp, vp, C.c_void_p, C.c_uint64, C.POINTER(vp)]),
        "ggq_comm_enable_peer": (C.c_int, [vp, vp, C.c_uint64, C.c_uint64]),
        "ggq_comm_peer_active": (C.c_int, [vp]),
        "ggq_get_rows_sharded_device": (C.c_int, [vp, vp, vp, vp, C.c_uint64, C.c_int, C.POINTER(vp), C.POINTER(vp), u64p]),
        "ggq_get_rows_sharded_slice_device": (C.c_int, [vp, vp, vp, vp, C.c_uint64, C.c_int, u64p, u64p, C.POINTER(vp), C.POINTER(vp), u64p]),
        "ggq_brwt_build_columns": (C.c_int, [u64p, C.c_uint64, C.c_uint64, C.c_int, C.c_uint32, C.c_uint64, C.c_int, C.POINTER(vp)]),
        "ggq_brwt_synthesize_greedy": (C.c_int, [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(vp)]),
        "ggq_tree_dfs_columns": (C.c_int, [vp, C.POINTER(C.c_uint32)]),
        "ggq_tree_shape": (C.c_int, [vp, u64p, C.POINTER(C.c_int64), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), u64p, u64p, C.POINTER(C.c_uint32)]),
        "ggq_brwt_serialize": (C.c_int, [vp, C.c_void_p, vp, u64p]),
        "ggq_brwt_serialize_buffer": (C.c_int, [vp, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_size_t)]),
        "ggq_rrr63_encode_buffer": (C.c_int, [u64p, C.c_uint64, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_size_t)]),
        "ggq_brwt_inspect_buffer": (C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t), u64p]),
        "ggq_brwt_inspect_columns": (C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
        "ggq_num_rows": (C.c_uint64, [vp]),
        "ggq_num_columns": (C.c_uint64, [vp]),
        "ggq_num_relations": (C.c_uint64, [vp]),
        "ggq_num_nodes": (C.c_uint64, [vp]),
        "ggq_total_column_size": (C.c_uint64, [vp]),
        "ggq_device_bytes": (C.c_uint64, [vp]),
        "ggq_slot_bits": (C.c_uint32, [vp]),
        "ggq_rrr63_decode_buffer": (C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t), u64p, C.c_uint64, u64p]),
        "ggq_brwt_serialize_file": (C.c_int, [vp, C.c_char_p, u64p]),
        "ggq_brwt_load_file": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(vp)]),
        "ggq_set_default_option": (C.c_int, [C.c_char_p, C.c_uint64]),
        "ggq_leaf_code_info": (C.c_int, [vp, C.POINTER(C.c_uint64)]),
        "ggq_last_probe_count": (C.c_uint64, [vp]),
        "ggq_set_option": (C.c_int, [vp, C.c_char_p, C.c_uint64]),
        "ggq_get_rows": (C.c_int, [vp, u64p, C.c_uint64, C.POINTER(vp)]),
        "ggq_rows_offsets": (u64p, [vp]),
        "ggq_rows_columns": (u64p, [vp]),
        "ggq_rows_count": (C.c_uint64, [vp]),
        "ggq_rows_result_free": (None, [vp]),
        "ggq_get_rows_into": (C.c_int, [vp, u64p, C.c_uint64, u64p, u64p, C.c_uint64, u64p]),
        "ggq_get_column_count": (C.c_int, [vp, C.c_uint64, u64p]),
        "ggq_get_column_into": (C.c_int, [vp, C.c_uint64, u64p, C.c_uint64, u64p]),
        "ggq_get_column_noalloc": (C.c_int, [vp, C.c_uint64, u64p, C.c_uint64, u64p]),
        "ggq_image_save": (C.c_int, [vp, C.c_char_p]),
        "ggq_image_load": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(vp)]),
        "ggq_get_labels": (C.c_int, [vp, u64p, C.c_uint64, C.c_double, u64p, u64p, u64p]),
        "ggq_get_top_labels": (C.c_int, [vp, u64p, C.c_uint64, C.c_uint64, u64p, u64p, u64p]),
        "ggq_get_labels_batch": (C.c_int, [vp, u64p, u64p, C.c_uint64, C.c_double, u64p, u64p, u64p, C.c_uint64, u64p]),
        "ggq_get_labels_batch_ratios": (C.c_int, [vp, u64p, u64p, C.c_uint64, C.POINTER(C.c_double), u64p, u64p, u64p, C.c_uint64, u64p]),
        "ggq_get_rows_into_u32": (C.c_int, [vp, u64p, C.c_uint64, u64p, C.c_void_p, C.c_uint64, u64p]),
        "ggq_get_rows_compact": (C.c_int, [vp, C.c_void_p, C.c_int, C.c_uint64, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_uint64, u64p]),
        "ggq_count_labels": (C.c_int, [vp, u64p, C.c_uint64, u64p]),
        "ggq_count_labels_device": (C.c_int, [vp, vp, vp, C.c_uint64, C.POINTER(vp)]),
        "ggq_slice_rows": (C.c_int, [vp, u64p, C.c_uint64, C.POINTER(u64p), u64p]),
        "ggq_buffer_free": (None, [vp]),
        "ggq_workspace_create": (C.c_int, [vp, C.POINTER(vp)]),
        "ggq_workspace_free": (None, [vp]),
        "ggq_workspace_stream": (vp, [vp]),
        "ggq_workspace_stat": (C.c_double, [vp, C.c_int]),
        "ggq_get_rows_device": (C.c_int, [vp, vp, vp, C.c_uint64, C.c_int, C.POINTER(vp), C.POINTER(vp), u64p]),
        "ggq_get": (C.c_int, [vp, u64p, u64p, C.c_uint64, u8p]),
        "ggq_has_column": (C.c_int, [vp, u64p, C.c_uint64, C.c_uint64, u8p]),
        "ggq_get_column": (C.c_int, [vp, C.c_uint64, C.POINTER(u64p), u64p]),
        "ggq_bv_load_buffer": (C.c_int, [C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t), C.c_int, C.POINTER(vp)]),
        "ggq_bv_build": (C.c_int, [C.c_int, u64p, C.c_uint64, C.c_int, C.POINTER(vp)]),
        "ggq_bv_synthesize": (C.c_int, [C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(vp)]),
        "ggq_bv_free": (None, [vp]),
        "ggq_bv_size": (C.c_uint64, [vp]),
        "ggq_bv_num_set_bits": (C.c_uint64, [vp]),
        "ggq_bv_is_inverted": (C.c_int, [vp]),
        "ggq_bv_device_bytes": (C.c_uint64, [vp]),
        "ggq_bv_rank1": (C.c_int, [vp, u64p, C.c_uint64, u64p]),
        "ggq_bv_select1": (C.c_int, [vp, u64p, C.c_uint64, u64p]),
        "ggq_bv_get": (C.c_int, [vp, u64p, C.c_uint64, u8p]),
        "ggq_bv_rank1_device": (C.c_int, [vp, vp, C.c_uint64, vp, vp]),
        "ggq_bv_select1_device": (C.c_int, [vp, vp, C.c_uint64, vp, vp]),
        "ggq_bv_get_device": (C.c_int, [vp, vp, C.c_uint64, vp, vp]),
        "ggq_map_rank1": (C.c_int, [vp, u64p, C.c_uint64, u64p, u64p]),
        "ggq_map_rank1_device": (C.c_int, [vp, vp, C.c_uint64, vp, vp]),
        "ggq_bv_rank0": (C.c_int, [vp, u64p, C.c_uint64, u64p]),
        "ggq_bv_select0": (C.c_int, [vp, u64p, C.c_uint64, u64p]),
        "ggq_bv_rank0_device": (C.c_int, [vp, vp, C.c_uint64, vp, vp]),
        "ggq_bv_select0_device": (C.c_int, [vp, vp, C.c_uint64, vp, vp]),
        "ggq_bv_get_int": (C.c_int, [vp, u64p, C.c_uint64, C.c_uint32, u64p]),
        "ggq_bv_to_words": (C.c_int, [vp, u64p]),
        "ggq_bv_call_ones_into": (C.c_int, [vp, u64p, C.c_uint64, u64p]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    L._ggq_signatures = sig
    _lib = L
    return L


def _check(rc):
    if rc != GGQ_OK:
        raise GgqError(rc, lib().ggq_last_error().decode("utf-8", "replace"))


def _u64(a):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    return a, a.ctypes.data_as(u64p)


NOALLOC_MAX_COLUMN_ROWS = 1 << 20   # GGQ_NOALLOC_MAX_COLUMN_ROWS (include/ggq.h)


def set_default_option(name, value):
    """Process-wide default for matrices created afterwards (ggq_set_default_option): "wide_slots" = 1 selects the
    capacity layout (32-byte slots of 224 positions)."""
    _check(lib().ggq_set_default_option(name.encode(), int(value)))


def device_count():
    n = C.c_int()
    rc = lib().ggq_device_count(C.byref(n))
    return n.value if rc == GGQ_OK else 0


def mix64(x):
    """splitmix64 finaliser on numpy uint64 arrays — the portable generator of SURVEY.md §8(d)."""
    x = np.asarray(x, dtype=np.uint64).copy()
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(30)
        x *= np.uint64(0xBF58476D1CE4E5B9)
        x ^= x >> np.uint64(27)
        x *= np.uint64(0x94D049BB133111EB)
        x ^= x >> np.uint64(31)
    return x


def density_threshold(d):
    """T = floor(d * 2^64), computed once in extended precision."""
    t = int(np.floor(np.longdouble(d) * np.longdouble(2.0) ** 64))
    return min(max(t, 0), (1 << 64) - 1)


SEED_MUL = 0x9E3779B97F4A7C15


def shard_plan(data, n_shards):
    """Host-only: (col_owner[num_columns], cut_depth) of the column-sharding plan for a serialized BRWT."""
    st = inspect_brwt(data)
    owner = np.zeros(max(int(st["num_columns"]), 1), np.uint32)
    cut = C.c_uint32()
    _check(lib().ggq_brwt_shard_plan_buffer(data, len(data), n_shards, owner.ctypes.data_as(C.POINTER(C.c_uint32)),
                                            C.byref(cut)))
    return owner[:int(st["num_columns"])], cut.value


class Comm:
    """NCCL communicator over the shards of one matrix (ggq_comm_*).  `id_bytes`: 128 bytes made by rank 0 with
    Comm.unique_id() and shipped to the other ranks (e.g. torch.distributed.broadcast_object_list)."""

    @staticmethod
    def unique_id():
        buf = (C.c_uint8 * 128)()
        _check(lib().ggq_comm_unique_id(buf))
        return bytes(buf)

    def __init__(self, id_bytes, rank, n_ranks, device):
        buf = (C.c_uint8 * 128).from_buffer_copy(id_bytes)
        h = C.c_void_p()
        _check(lib().ggq_comm_init(buf, rank, n_ranks, device, C.byref(h)))
        self._h = h

    def enable_peer(self, workspace, max_rows, max_ids):
        """Collective: export `workspace`'s partial-CSR buffers over cudaIpc and map the peers' (ggq_comm_enable_peer);
        sharded queries on that workspace then merge straight from peer memory over NVLink."""
        _check(lib().ggq_comm_enable_peer(self._h, workspace._h, int(max_rows), int(max_ids)))

    @property
    def peer_active(self):
        return bool(lib().ggq_comm_peer_active(self._h))

    def close(self):
        if self._h:
            lib().ggq_comm_free(self._h)
            self._h = None


def rrr63_encode(words, n_bits):
    """Host only: bytes of sdsl's rrr_vector<63>(bits).serialize() for the given plain bits."""
    w = np.ascontiguousarray(words, dtype=np.uint64)
    if w.size == 0:
        w = np.zeros(1, np.uint64)
    out = C.POINTER(C.c_uint8)()
    n = C.c_size_t()
    _check(lib().ggq_rrr63_encode_buffer(w.ctypes.data_as(u64p), n_bits, C.byref(out), C.byref(n)))
    data = C.string_at(out, n.value)
    lib().ggq_buffer_free(out)
    return data


def rrr63_decode(data):
    """Host only: (plain bits as LSB-first uint64 words, n_bits, bytes consumed) of one serialized rrr_vector<63>, by the
    loader's own decoder (ggq_rrr63_decode_buffer)."""
    n_bits, used = C.c_uint64(), C.c_size_t()
    rc = lib().ggq_rrr63_decode_buffer(data, len(data), C.byref(used), None, 0, C.byref(n_bits))
    if rc != 0 and n_bits.value == 0:
        _check(rc)
    w = np.zeros(max((n_bits.value + 63) // 64, 1), np.uint64)
    _check(lib().ggq_rrr63_decode_buffer(data, len(data), C.byref(used), w.ctypes.data_as(u64p), w.size, C.byref(n_bits)))
    return w[: (n_bits.value + 63) // 64], n_bits.value, used.value


def inspect_brwt(data):
    """Host-only parse + flatten of a serialized BRWT (no GPU needed) -> dict of stats."""
    stats = (C.c_uint64 * 8)()
    used = C.c_size_t()
    _check(lib().ggq_brwt_inspect_buffer(data, len(data), C.byref(used), stats))
    keys = ["num_rows", "num_columns", "num_nodes", "num_relations", "total_column_size", "groups",
            "device_bytes", "max_depth"]
    d = dict(zip(keys, [int(x) for x in stats]))
    d["consumed"] = used.value
    m = d["num_columns"]
    dfs = np.zeros(max(m, 1), dtype=np.uint32)
    dep = np.zeros(max(m, 1), dtype=np.uint32)
    _check(lib().ggq_brwt_inspect_columns(data, len(data), dfs.ctypes.data_as(C.POINTER(C.c_uint32)),
                                          dep.ctypes.data_as(C.POINTER(C.c_uint32))))
    d["dfs_columns"] = dfs[:m].copy()
    d["column_depth"] = dep[:m].copy()
    return d


class GpuMatrix:
    """GPU-resident BinaryMatrix (reference interface: common/binary_matrix.hpp:9-29).

    ``get_row(i)`` returns exactly what the reference's ``get_row(i)`` returns (same order);
    ``get_rows / slice_rows / has_column`` are the batched compositions named by the north star.
    """

    def __init__(self, handle, kind):
        self._h = C.c_void_p(handle)
        self.kind = kind

    # -- construction -----------------------------------------------------------------------
    @classmethod
    def load_brwt(cls, data, device=0):
        """BRWT::load on the reference's serialized stream.  Returns (matrix, bytes_consumed)."""
        h = C.c_void_p()
        used = C.c_size_t()
        _check(lib().ggq_brwt_load_buffer(data, len(data), C.byref(used), device, C.byref(h)))
        m = cls(h.value, "brwt")
        m.consumed = used.value
        return m

    @classmethod
    def load_brwt_file(cls, path, device=0):
        """BRWT::load from a file holding the matrix stream (ggq_brwt_load_file)."""
        h = C.c_void_p()
        _check(lib().ggq_brwt_load_file(os.fsencode(path), device, C.byref(h)))
        return cls(h.value, "brwt")

    def serialize_to_file(self, path):
        """BRWT::serialize from the device image into a file (ggq_brwt_serialize_file) -> bytes written."""
        n = C.c_uint64()
        _check(lib().ggq_brwt_serialize_file(self._h, os.fsencode(path), C.byref(n)))
        return n.value

    @classmethod
    def build_brwt(cls, col_words, n_rows, greedy=False, arity=2, relax=0, device=0):
        """The reference's bottom-up construction on the device (ggq_brwt_build_columns): col_words[j] = column j as
        ceil(n_rows / 64) LSB-first u64 words."""
        cw = np.ascontiguousarray(col_words, dtype=np.uint64)
        m = cw.shape[0] if cw.ndim == 2 else 0
        if cw.size == 0:
            cw = np.zeros(1, np.uint64)
        h = C.c_void_p()
        _check(lib().ggq_brwt_build_columns(cw.ctypes.data_as(u64p), n_rows, m, int(greedy), arity, relax, device, C.byref(h)))
        return cls(h.value, "brwt")

    def serialize(self):
        """BRWT::serialize from the device image (ggq_brwt_serialize_buffer) -> bytes."""
        out = C.POINTER(C.c_uint8)()
        n = C.c_size_t()
        _check(lib().ggq_brwt_serialize_buffer(self._h, C.byref(out), C.byref(n)))
        data = C.string_at(out, n.value)
        lib().ggq_buffer_free(out)
        return data

    @classmethod
    def synthesize_brwt_greedy(cls, n_rows, n_cols, density, seed=42, relax=7, device=0):
        """The implicit matrix of synthesize_brwt under the reference's greedy partitioner + relax(relax)."""
        h = C.c_void_p()
        _check(lib().ggq_brwt_synthesize_greedy(n_rows, n_cols, density_threshold(density), seed, relax, device, C.byref(h)))
        return cls(h.value, "brwt")

    def dfs_columns(self):
        """Columns in the order get_row lists them (DFS leaf order)."""
        out = np.zeros(max(int(self.num_columns), 1), np.uint32)
        _check(lib().ggq_tree_dfs_columns(self._h, out.ctypes.data_as(C.POINTER(C.c_uint32))))
        return out[:int(self.num_columns)]

    def tree_shape(self):
        """dict of per-node arrays in level order: parent, n_children, n_columns, size, ones, first_column."""
        n = C.c_uint64()
        _check(lib().ggq_tree_shape(self._h, C.byref(n), None, None, None, None, None, None))
        k = n.value
        out = {"parent": np.zeros(k, np.int64), "n_children": np.zeros(k, np.uint32), "n_columns": np.zeros(k, np.uint32),
               "size": np.zeros(k, np.uint64), "ones": np.zeros(k, np.uint64), "first_column": np.zeros(k, np.uint32)}
        _check(lib().ggq_tree_shape(self._h, C.byref(n), out["parent"].ctypes.data_as(C.POINTER(C.c_int64)),
                                    out["n_children"].ctypes.data_as(C.POINTER(C.c_uint32)),
                                    out["n_columns"].ctypes.data_as(C.POINTER(C.c_uint32)),
                                    out["size"].ctypes.data_as(u64p), out["ones"].ctypes.data_as(u64p),
                                    out["first_column"].ctypes.data_as(C.POINTER(C.c_uint32))))
        return out

    @classmethod
    def load_image(cls, path, device=0):
        """Flattened device image written by save_image (ggq_image_load): uploaded as it is."""
        h = C.c_void_p()
        _check(lib().ggq_image_load(os.fsencode(path), device, C.byref(h)))
        return cls(h.value, "brwt")

    def save_image(self, path):
        _check(lib().ggq_image_save(self._h, os.fsencode(path)))

    @classmethod
    def load_brwt_shard(cls, data, shard, n_shards, device=0):
        """Column shard `shard` of `n_shards` of a serialized BRWT (ggq_brwt_load_shard_buffer)."""
        h = C.c_void_p()
        used = C.c_size_t()
        _check(lib().ggq_brwt_load_shard_buffer(data, len(data), C.byref(used), device, shard, n_shards, C.byref(h)))
        m = cls(h.value, "brwt")
        m.consumed = used.value
        return m

    @classmethod
    def synthesize_brwt_shard(cls, n_rows, n_cols, density, shard, n_shards, seed=42, arity=4, device=0):
        h = C.c_void_p()
        t = density_threshold(density)
        _check(lib().ggq_brwt_synthesize_shard(n_rows, n_cols, t, seed, arity, device, shard, n_shards, C.byref(h)))
        m = cls(h.value, "brwt")
        m.threshold, m.seed = t, seed
        return m

    def shard_info(self):
        """-> (shard, n_shards, col_owned[num_columns] as bool array)."""
        a, b = C.c_uint32(), C.c_uint32()
        own = np.zeros(max(int(self.num_columns), 1), np.uint8)
        _check(lib().ggq_shard_info(self._h, C.byref(a), C.byref(b), own.ctypes.data_as(u8p)))
        return a.value, b.value, own[:int(self.num_columns)].astype(bool)

    @classmethod
    def load_colmajor(cls, data, device=0):
        h = C.c_void_p()
        used = C.c_size_t()
        _check(lib().ggq_colmajor_load_buffer(data, len(data), C.byref(used), device, C.byref(h)))
        m = cls(h.value, "colmajor")
        m.consumed = used.value
        return m

    @classmethod
    def synthesize_brwt(cls, n_rows, n_cols, density, seed=42, arity=4, device=0, threshold=None):
        h = C.c_void_p()
        t = density_threshold(density) if threshold is None else threshold
        _check(lib().ggq_brwt_synthesize(n_rows, n_cols, t, seed, arity, device, C.byref(h)))
        m = cls(h.value, "brwt")
        m.threshold, m.seed = t, seed
        return m

    def close(self):
        if self._h:
            lib().ggq_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- scalars ----------------------------------------------------------------------------
    num_rows = property(lambda s: lib().ggq_num_rows(s._h))
    num_columns = property(lambda s: lib().ggq_num_columns(s._h))
    num_relations = property(lambda s: lib().ggq_num_relations(s._h))
    num_nodes = property(lambda s: lib().ggq_num_nodes(s._h))
    total_column_size = property(lambda s: lib().ggq_total_column_size(s._h))
    device_bytes = property(lambda s: lib().ggq_device_bytes(s._h))
    slot_bits = property(lambda s: lib().ggq_slot_bits(s._h))

    @property
    def leaf_codes(self):
        """Leaf codes of a BRWT image: dict(nodes, bytes, escape_rate, in_use)."""
        w = (C.c_uint64 * 4)()
        _check(lib().ggq_leaf_code_info(self._h, w))
        return {"nodes": int(w[0]), "bytes": int(w[1]), "escape_rate": w[2] / 1e6, "in_use": bool(w[3])}
    last_probe_count = property(lambda s: lib().ggq_last_probe_count(s._h))

    def set_option(self, name, value):
        """Tuning knob of the row-query path (ggq_set_option): rows_per_warp, sort_bits, sort_min_rows."""
        _check(lib().ggq_set_option(self._h, name.encode(), int(value)))

    # -- queries (host buffers) ---------------------------------------------------------------
    def get_rows_csr(self, rows):
        """-> (offsets[n+1], cols[offsets[n]]) as uint64 arrays."""
        rows, rp = _u64(rows)
        res = C.c_void_p()
        _check(lib().ggq_get_rows(self._h, rp, rows.size, C.byref(res)))
        try:
            n = rows.size
            off = np.ctypeslib.as_array(lib().ggq_rows_offsets(res), shape=(n + 1,)).copy()
            total = int(off[-1])
            if total:
                cols = np.ctypeslib.as_array(lib().ggq_rows_columns(res), shape=(total,)).copy()
            else:
                cols = np.zeros(0, dtype=np.uint64)
        finally:
            lib().ggq_rows_result_free(res)
        return off, cols

    def get_rows_into(self, rows, offsets, cols):
        """Caller-provided (e.g. pinned) output arrays; returns the number of column ids.  `cols` may be uint64
        (ggq_get_rows_into) or uint32 (ggq_get_rows_into_u32, half the device->host bytes)."""
        n_cols = C.c_uint64()
        if cols.dtype == np.uint32:
            _check(lib().ggq_get_rows_into_u32(self._h, rows.ctypes.data_as(u64p), rows.size,
                                               offsets.ctypes.data_as(u64p), cols.ctypes.data, cols.size,
                                               C.byref(n_cols)))
        else:
            assert cols.dtype == np.uint64
            _check(lib().ggq_get_rows_into(self._h, rows.ctypes.data_as(u64p), rows.size,
                                           offsets.ctypes.data_as(u64p), cols.ctypes.data_as(u64p),
                                           cols.size, C.byref(n_cols)))
        return n_cols.value

    def get_rows_compact(self, rows, counts, cols):
        """ggq_get_rows_compact: rows u32 or u64, counts u8 / u16 / u32 (one per row), cols u16 / u32 / u64 — all
        caller-provided numpy arrays (pinned for speed).  Returns the number of column ids."""
        assert rows.dtype in (np.uint32, np.uint64) and counts.dtype in (np.uint8, np.uint16, np.uint32)
        assert cols.dtype in (np.uint16, np.uint32, np.uint64) and counts.size >= rows.size
        n = C.c_uint64()
        _check(lib().ggq_get_rows_compact(self._h, rows.ctypes.data, rows.dtype.itemsize, rows.size, counts.ctypes.data,
                                          counts.dtype.itemsize, cols.ctypes.data, cols.dtype.itemsize, cols.size, C.byref(n)))
        return n.value

    def get_rows(self, rows):
        off, cols = self.get_rows_csr(rows)
        return [cols[off[i]:off[i + 1]] for i in range(len(off) - 1)]

    def get_row(self, row):
        return self.get_rows_csr([row])[1]

    def count_labels(self, rows):
        """Histogram of set columns over the multiset `rows` (ggq_count_labels) -> uint64[num_columns]."""
        rows, rp = _u64(rows)
        out = np.zeros(max(int(self.num_columns), 1), np.uint64)
        _check(lib().ggq_count_labels(self._h, rp, rows.size, out.ctypes.data_as(u64p)))
        return out[:int(self.num_columns)]

    def get_labels(self, rows, presence_ratio=0.0):
        """Columns present in at least ceil(len(rows) * ratio) of the rows (ratio 0: in any), ascending, with
        their counts (ggq_get_labels; StaticBinRelAnnotator::get_labels(indices, ratio))."""
        rows, rp = _u64(rows)
        m = max(int(self.num_columns), 1)
        cols, cnt, k = np.zeros(m, np.uint64), np.zeros(m, np.uint64), C.c_uint64()
        _check(lib().ggq_get_labels(self._h, rp, rows.size, float(presence_ratio), cols.ctypes.data_as(u64p),
                                    cnt.ctypes.data_as(u64p), C.byref(k)))
        return cols[:k.value], cnt[:k.value]

    def get_labels_batch(self, rows, query_offsets, presence_ratio=0.0, capacity=None):
        """get_labels for many queries in one launch (ggq_get_labels_batch / _ratios): query q owns
        rows[query_offsets[q]:query_offsets[q+1]]; presence_ratio is one number or one per query.
        Returns (label_offsets, cols, counts)."""
        rows, rp = _u64(rows)
        qo, qp = _u64(query_offsets)
        nq = qo.size - 1
        if capacity is None:
            capacity = max(nq * int(self.num_columns), 1)
        loff = np.zeros(nq + 1, np.uint64)
        cols, cnt, k = np.zeros(capacity, np.uint64), np.zeros(capacity, np.uint64), C.c_uint64()
        if np.ndim(presence_ratio) == 0:
            _check(lib().ggq_get_labels_batch(self._h, rp, qp, nq, float(presence_ratio), loff.ctypes.data_as(u64p),
                                              cols.ctypes.data_as(u64p), cnt.ctypes.data_as(u64p), capacity, C.byref(k)))
        else:
            pr = np.ascontiguousarray(presence_ratio, dtype=np.float64)
            assert pr.size == nq
            _check(lib().ggq_get_labels_batch_ratios(self._h, rp, qp, nq, pr.ctypes.data_as(C.POINTER(C.c_double)),
                                                     loff.ctypes.data_as(u64p), cols.ctypes.data_as(u64p),
                                                     cnt.ctypes.data_as(u64p), capacity, C.byref(k)))
        return loff, cols[:k.value], cnt[:k.value]

    def get_top_labels(self, rows, num_top):
        """The num_top most frequent columns over the multiset `rows`, decreasing count, ties by ascending
        column (ggq_get_top_labels; MultiLabelEncoded::get_top_labels)."""
        rows, rp = _u64(rows)
        m = max(min(int(self.num_columns), int(num_top)), 1)
        cols, cnt, k = np.zeros(m, np.uint64), np.zeros(m, np.uint64), C.c_uint64()
        _check(lib().ggq_get_top_labels(self._h, rp, rows.size, int(num_top), cols.ctypes.data_as(u64p),
                                        cnt.ctypes.data_as(u64p), C.byref(k)))
        return cols[:k.value], cnt[:k.value]

    def slice_rows(self, rows):
        rows, rp = _u64(rows)
        out = u64p()
        n = C.c_uint64()
        _check(lib().ggq_slice_rows(self._h, rp, rows.size, C.byref(out), C.byref(n)))
        try:
            res = np.ctypeslib.as_array(out, shape=(n.value,)).copy() if n.value else np.zeros(0, np.uint64)
        finally:
            lib().ggq_buffer_free(out)
        return res

    def get_cells(self, rows, cols):
        rows, rp = _u64(rows)
        cols, cp = _u64(cols)
        out = np.empty(rows.size, dtype=np.uint8)
        _check(lib().ggq_get(self._h, rp, cp, rows.size, out.ctypes.data_as(u8p)))
        return out

    def get(self, row, col):
        return bool(self.get_cells([row], [col])[0])

    def has_column(self, rows, col):
        rows, rp = _u64(rows)
        out = np.empty(rows.size, dtype=np.uint8)
        _check(lib().ggq_has_column(self._h, rp, rows.size, col, out.ctypes.data_as(u8p)))
        return out

    def get_column_count(self, col):
        n = C.c_uint64()
        _check(lib().ggq_get_column_count(self._h, col, C.byref(n)))
        return n.value

    def get_column_into(self, col, out):
        """Ascending set rows of `col` into the caller's uint64 array (e.g. pinned); returns their number."""
        n = C.c_uint64()
        _check(lib().ggq_get_column_into(self._h, col, out.ctypes.data_as(u64p), out.size, C.byref(n)))
        return n.value

    def get_column_noalloc(self, col, out):
        """get_column_into through the entry point that allocates nothing on the device (ggq_get_column_noalloc: fixed
        scratch inside the library, at most NOALLOC_MAX_COLUMN_ROWS set rows); returns their number."""
        n = C.c_uint64()
        _check(lib().ggq_get_column_noalloc(self._h, col, out.ctypes.data_as(u64p), out.size, C.byref(n)))
        return n.value

    def get_column(self, col):
        out = u64p()
        n = C.c_uint64()
        _check(lib().ggq_get_column(self._h, col, C.byref(out), C.byref(n)))
        try:
            res = np.ctypeslib.as_array(out, shape=(n.value,)).copy() if n.value else np.zeros(0, np.uint64)
        finally:
            lib().ggq_buffer_free(out)
        return res

    # -- device-resident API (bench.py) -----------------------------------------------------
    def workspace(self):
        return Workspace(self)


class Workspace:
    def __init__(self, matrix):
        self.m = matrix
        h = C.c_void_p()
        _check(lib().ggq_workspace_create(matrix._h, C.byref(h)))
        self._h = h

    @property
    def stream(self):
        return lib().ggq_workspace_stream(self._h)

    def stat(self, what):
        """what: 'kernel_ms' | 'probes' | 'touches' | 'launches' of the last query on this workspace."""
        return lib().ggq_workspace_stat(self._h, {"kernel_ms": 0, "probes": 1, "touches": 2, "launches": 3,
                                                  "exchange_ms": 4, "exchange_bytes": 5, "code_probes": 6,
                                                  "code_escapes": 7}[what])

    def get_rows_device(self, d_rows_ptr, n, col_bytes=4):
        """d_rows_ptr: device pointer (int).  -> (d_offsets_ptr, d_cols_ptr, n_cols), owned by the workspace."""
        off = C.c_void_p()
        cols = C.c_void_p()
        total = C.c_uint64()
        _check(lib().ggq_get_rows_device(self.m._h, self._h, C.c_void_p(d_rows_ptr), n, col_bytes,
                                         C.byref(off), C.byref(cols), C.byref(total)))
        return off.value, cols.value, total.value

    def count_labels_device(self, d_rows_ptr, n):
        """-> device pointer to uint64[num_columns] counts, owned by the workspace."""
        out = C.c_void_p()
        _check(lib().ggq_count_labels_device(self.m._h, self._h, C.c_void_p(d_rows_ptr), n, C.byref(out)))
        return out.value

    def count_labels_allreduce_device(self, comm, d_rows_ptr, n):
        """Collective: histogram of the whole batch on every rank (ggq_count_labels_allreduce_device) -> device ptr."""
        out = C.c_void_p()
        _check(lib().ggq_count_labels_allreduce_device(self.m._h, self._h, comm._h, C.c_void_p(d_rows_ptr), n,
                                                       C.byref(out)))
        return out.value

    def get_rows_sharded_device(self, comm, d_rows_ptr, n, col_bytes=4):
        """Collective over all shards (same batch on every rank) -> complete CSR on every rank."""
        off, cols, total = C.c_void_p(), C.c_void_p(), C.c_uint64()
        _check(lib().ggq_get_rows_sharded_device(self.m._h, self._h, comm._h, C.c_void_p(d_rows_ptr), n, col_bytes,
                                                 C.byref(off), C.byref(cols), C.byref(total)))
        return off.value, cols.value, total.value

    def get_rows_sharded_slice_device(self, comm, d_rows_ptr, n, col_bytes=4):
        """Collective; this rank merges only its slice of the batch -> (first_query, n_queries, d_offsets, d_cols, n_cols)."""
        off, cols, total, first, cnt = C.c_void_p(), C.c_void_p(), C.c_uint64(), C.c_uint64(), C.c_uint64()
        _check(lib().ggq_get_rows_sharded_slice_device(self.m._h, self._h, comm._h, C.c_void_p(d_rows_ptr), n, col_bytes,
                                                       C.byref(first), C.byref(cnt), C.byref(off), C.byref(cols), C.byref(total)))
        return first.value, cnt.value, off.value, cols.value, total.value

    def merge_shards_device(self, n, d_offsets_ptrs, d_cols_ptrs, col_bytes=4):
        """Merge step alone on device pointers of every shard's partial CSR -> (d_offsets, d_cols, n_cols)."""
        R = len(d_offsets_ptrs)
        po = (C.c_void_p * R)(*d_offsets_ptrs)
        pc = (C.c_void_p * R)(*d_cols_ptrs)
        off, cols, total = C.c_void_p(), C.c_void_p(), C.c_uint64()
        _check(lib().ggq_merge_shards_device(self._h, n, R, po, pc, col_bytes, C.byref(off), C.byref(cols), C.byref(total)))
        return off.value, cols.value, total.value

    def close(self):
        if self._h:
            lib().ggq_workspace_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


BV_KINDS = {"stat": 0, "sd": 1, "rrr": 2}


class GpuBitVector:
    """GPU-resident bit_vector_{stat,sd,rrr} (reference interface: common/bit_vector.hpp:12-46)."""

    def __init__(self, handle, kind):
        self._h = C.c_void_p(handle)
        self.kind = kind

    @classmethod
    def build(cls, kind, words, n_bits, device=0):
        words = np.ascontiguousarray(words, dtype=np.uint64)
        if words.size == 0:
            words = np.zeros(1, dtype=np.uint64)
        h = C.c_void_p()
        _check(lib().ggq_bv_build(BV_KINDS[kind], words.ctypes.data_as(u64p), n_bits, device, C.byref(h)))
        return cls(h.value, kind)

    @classmethod
    def load(cls, kind, data, device=0):
        h = C.c_void_p()
        used = C.c_size_t()
        _check(lib().ggq_bv_load_buffer(BV_KINDS[kind], data, len(data), C.byref(used), device, C.byref(h)))
        v = cls(h.value, kind)
        v.consumed = used.value
        return v

    @classmethod
    def synthesize(cls, kind, n_bits, density, seed=42, device=0):
        h = C.c_void_p()
        _check(lib().ggq_bv_synthesize(BV_KINDS[kind], n_bits, density_threshold(density), seed, device, C.byref(h)))
        return cls(h.value, kind)

    def close(self):
        if self._h:
            lib().ggq_bv_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    size = property(lambda s: lib().ggq_bv_size(s._h))
    num_set_bits = property(lambda s: lib().ggq_bv_num_set_bits(s._h))
    is_inverted = property(lambda s: bool(lib().ggq_bv_is_inverted(s._h)))
    device_bytes = property(lambda s: lib().ggq_bv_device_bytes(s._h))

    def rank1(self, q):
        q, qp = _u64(q)
        out = np.empty(q.size, dtype=np.uint64)
        _check(lib().ggq_bv_rank1(self._h, qp, q.size, out.ctypes.data_as(u64p)))
        return out

    def select1(self, q):
        q, qp = _u64(q)
        out = np.empty(q.size, dtype=np.uint64)
        _check(lib().ggq_bv_select1(self._h, qp, q.size, out.ctypes.data_as(u64p)))
        return out

    def get(self, q):
        q, qp = _u64(q)
        out = np.empty(q.size, dtype=np.uint8)
        _check(lib().ggq_bv_get(self._h, qp, q.size, out.ctypes.data_as(u8p)))
        return out

    def map_rank1(self, q):
        """graph index -> annotation row (AnnotatedDBG::graph_to_anno_index): rank1(i) - 1, UINT64_MAX where unset.
        Returns (anno_idx, n_unmapped)."""
        q, qp = _u64(q)
        out = np.empty(q.size, dtype=np.uint64)
        bad = C.c_uint64()
        _check(lib().ggq_map_rank1(self._h, qp, q.size, out.ctypes.data_as(u64p), C.byref(bad)))
        return out, bad.value

    def rank0(self, q):
        q, qp = _u64(q)
        out = np.empty(q.size, dtype=np.uint64)
        _check(lib().ggq_bv_rank0(self._h, qp, q.size, out.ctypes.data_as(u64p)))
        return out

    def select0(self, q):
        q, qp = _u64(q)
        out = np.empty(q.size, dtype=np.uint64)
        _check(lib().ggq_bv_select0(self._h, qp, q.size, out.ctypes.data_as(u64p)))
        return out

    def get_int(self, q, width):
        q, qp = _u64(q)
        out = np.empty(q.size, dtype=np.uint64)
        _check(lib().ggq_bv_get_int(self._h, qp, q.size, width, out.ctypes.data_as(u64p)))
        return out

    def to_words(self):
        out = np.zeros(max((self.size + 63) // 64, 1), dtype=np.uint64)
        _check(lib().ggq_bv_to_words(self._h, out.ctypes.data_as(u64p)))
        return out[:(self.size + 63) // 64]

    def call_ones(self):
        n = C.c_uint64()
        out = np.empty(max(int(self.num_set_bits), 1), dtype=np.uint64)
        _check(lib().ggq_bv_call_ones_into(self._h, out.ctypes.data_as(u64p), out.size, C.byref(n)))
        return out[:n.value]

    def rank1_device(self, d_q, n, d_out, stream=0):
        _check(lib().ggq_bv_rank1_device(self._h, d_q, n, d_out, stream))

    def select1_device(self, d_q, n, d_out, stream=0):
        _check(lib().ggq_bv_select1_device(self._h, d_q, n, d_out, stream))

    def get_device(self, d_q, n, d_out, stream=0):
        _check(lib().ggq_bv_get_device(self._h, d_q, n, d_out, stream))
