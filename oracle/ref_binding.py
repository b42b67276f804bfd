"""TEST INFRASTRUCTURE — ctypes binding of oracle/_ref/libggref.so.

libggref.so is the UNMODIFIED reference (BRWT, ColMajorCompressed, bit_vector_*,
DataGenerator) behind the C glue in oracle/ref_capi.cpp; it is built by
oracle/Makefile from /root/reference.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference leg may import this module.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_ref", "libggref.so")

u64p = C.POINTER(C.c_uint64)
u8p = C.POINTER(C.c_uint8)


def available():
    return os.path.exists(_LIB_PATH)


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(_LIB_PATH)
    L.ggref_brwt_build.restype = C.c_void_p
    L.ggref_brwt_build.argtypes = [u64p, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_uint64, C.c_int]
    L.ggref_colmajor_build.restype = C.c_void_p
    L.ggref_colmajor_build.argtypes = [u64p, C.c_uint64, C.c_uint64]
    L.ggref_matrix_load.restype = C.c_void_p
    L.ggref_matrix_load.argtypes = [C.c_int, C.c_char_p, C.c_uint64]
    L.ggref_matrix_serialize.restype = C.c_void_p
    L.ggref_matrix_serialize.argtypes = [C.c_void_p, u64p]
    L.ggref_matrix_free.argtypes = [C.c_void_p]
    L.ggref_bytes_free.argtypes = [C.c_void_p]
    for f in ("num_rows", "num_columns", "num_relations"):
        fn = getattr(L, "ggref_matrix_" + f)
        fn.restype = C.c_uint64
        fn.argtypes = [C.c_void_p]
    L.ggref_brwt_num_nodes.restype = C.c_uint64
    L.ggref_brwt_num_nodes.argtypes = [C.c_void_p]
    L.ggref_brwt_total_column_size.restype = C.c_uint64
    L.ggref_brwt_total_column_size.argtypes = [C.c_void_p]
    L.ggref_brwt_avg_arity.restype = C.c_double
    L.ggref_brwt_avg_arity.argtypes = [C.c_void_p]
    L.ggref_matrix_get.restype = C.c_int
    L.ggref_matrix_get.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64]
    L.ggref_matrix_get_row.restype = C.c_uint64
    L.ggref_matrix_get_row.argtypes = [C.c_void_p, C.c_uint64, u64p, C.c_uint64]
    L.ggref_matrix_get_column.restype = C.c_void_p
    L.ggref_matrix_get_column.argtypes = [C.c_void_p, C.c_uint64, u64p]
    L.ggref_matrix_get_rows.restype = C.c_double
    L.ggref_matrix_get_rows.argtypes = [C.c_void_p, u64p, C.c_uint64, u64p, C.POINTER(C.c_void_p), C.c_int]
    L.ggref_matrix_time_get_rows.restype = C.c_double
    L.ggref_matrix_time_get_rows.argtypes = [C.c_void_p, u64p, C.c_uint64, C.c_int, u64p]
    L.ggref_matrix_time_get.restype = C.c_double
    L.ggref_matrix_time_get.argtypes = [C.c_void_p, u64p, u64p, C.c_uint64, C.c_int, u8p]
    L.ggref_bv_build.restype = C.c_void_p
    L.ggref_bv_build.argtypes = [C.c_int, u64p, C.c_uint64]
    L.ggref_bv_load.restype = C.c_void_p
    L.ggref_bv_load.argtypes = [C.c_int, C.c_char_p, C.c_uint64]
    L.ggref_bv_serialize.restype = C.c_void_p
    L.ggref_bv_serialize.argtypes = [C.c_void_p, u64p]
    L.ggref_bv_free.argtypes = [C.c_void_p]
    L.ggref_bv_size.restype = C.c_uint64
    L.ggref_bv_size.argtypes = [C.c_void_p]
    L.ggref_bv_num_set_bits.restype = C.c_uint64
    L.ggref_bv_num_set_bits.argtypes = [C.c_void_p]
    L.ggref_bv_is_inverted.restype = C.c_int
    L.ggref_bv_is_inverted.argtypes = [C.c_void_p]
    L.ggref_bv_query.restype = C.c_double
    L.ggref_bv_query.argtypes = [C.c_void_p, C.c_int, u64p, C.c_uint64, u64p, C.c_int]
    L.ggref_gen_norepl.argtypes = [C.c_uint64, C.c_uint64, C.c_double, C.c_int, u64p]
    L.ggref_gen_uniform_rows.argtypes = [C.c_uint64, C.c_uint64, C.c_double, C.c_uint64, C.c_int, u64p]
    L.ggref_gen_random_ints.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, u64p]
    L.ggref_gen_implicit.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, u64p]
    L.ggref_max_threads.restype = C.c_int
    _lib = L
    return L


def _p(a):
    return a.ctypes.data_as(u64p)


def _take_bytes(ptr, n):
    out = C.string_at(ptr, n)
    lib().ggref_bytes_free(ptr)
    return out


def pack_columns(dense):
    """dense: (n, m) bool/0-1 array -> column-major u64 words (m, ceil(n/64)), LSB-first."""
    dense = np.asarray(dense, dtype=np.uint8)
    n, m = dense.shape
    wpc = (n + 63) // 64
    padded = np.zeros((m, wpc * 64), dtype=np.uint8)
    padded[:, :n] = dense.T
    return np.packbits(padded, axis=1, bitorder="little").view(np.uint64).reshape(m, wpc).copy()


def unpack_columns(words, n):
    """inverse of pack_columns -> (n, m) uint8."""
    m = words.shape[0]
    bits = np.unpackbits(words.view(np.uint8).reshape(m, -1), axis=1, bitorder="little")
    return bits[:, :n].T.copy()


class RefMatrix:
    """A reference BinaryMatrix (BRWT or ColMajorCompressed)."""

    def __init__(self, handle):
        if not handle:
            raise ValueError("reference load()/build failed")
        self.h = C.c_void_p(handle)

    @classmethod
    def build_brwt(cls, col_words, n, arity=2, greedy=False, relax=0, threads=1):
        col_words = np.ascontiguousarray(col_words, dtype=np.uint64)
        m = col_words.shape[0] if n or col_words.size else 0
        if col_words.ndim == 2:
            m = col_words.shape[0]
        return cls(lib().ggref_brwt_build(_p(col_words), n, m, arity, int(greedy), relax, threads))

    @classmethod
    def build_colmajor(cls, col_words, n):
        col_words = np.ascontiguousarray(col_words, dtype=np.uint64)
        return cls(lib().ggref_colmajor_build(_p(col_words), n, col_words.shape[0]))

    @classmethod
    def load(cls, kind, data):
        """kind: 'brwt' | 'colmajor'.  Raises ValueError when the reference load() returns false."""
        return cls(lib().ggref_matrix_load(0 if kind == "brwt" else 1, data, len(data)))

    def serialize(self):
        n = C.c_uint64()
        ptr = lib().ggref_matrix_serialize(self.h, C.byref(n))
        return _take_bytes(ptr, n.value)

    def __del__(self):
        if getattr(self, "h", None):
            lib().ggref_matrix_free(self.h)
            self.h = None

    num_rows = property(lambda s: lib().ggref_matrix_num_rows(s.h))
    num_columns = property(lambda s: lib().ggref_matrix_num_columns(s.h))
    num_relations = property(lambda s: lib().ggref_matrix_num_relations(s.h))
    num_nodes = property(lambda s: lib().ggref_brwt_num_nodes(s.h))
    total_column_size = property(lambda s: lib().ggref_brwt_total_column_size(s.h))
    avg_arity = property(lambda s: lib().ggref_brwt_avg_arity(s.h))

    def get(self, row, col):
        return bool(lib().ggref_matrix_get(self.h, row, col))

    def get_row(self, row):
        cap = max(1, int(self.num_columns))
        out = np.empty(cap, dtype=np.uint64)
        k = lib().ggref_matrix_get_row(self.h, row, _p(out), cap)
        return out[:k].copy()

    def get_column(self, col):
        n = C.c_uint64()
        ptr = lib().ggref_matrix_get_column(self.h, col, C.byref(n))
        out = np.frombuffer(C.string_at(ptr, n.value * 8), dtype=np.uint64).copy()
        lib().ggref_bytes_free(ptr)
        return out

    def get_rows(self, rows, threads=1):
        """-> (offsets[n+1], cols[sum k], seconds)"""
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        offsets = np.empty(rows.size + 1, dtype=np.uint64)
        ptr = C.c_void_p()
        secs = lib().ggref_matrix_get_rows(self.h, _p(rows), rows.size, _p(offsets), C.byref(ptr), threads)
        total = int(offsets[-1])
        cols = np.frombuffer(C.string_at(ptr, total * 8), dtype=np.uint64).copy()
        lib().ggref_bytes_free(ptr)
        return offsets, cols, secs

    def time_get_rows(self, rows, threads):
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        chk = C.c_uint64()
        secs = lib().ggref_matrix_time_get_rows(self.h, _p(rows), rows.size, threads, C.byref(chk))
        return secs, chk.value

    def get_cells(self, rows, cols, threads=1):
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        cols = np.ascontiguousarray(cols, dtype=np.uint64)
        out = np.empty(rows.size, dtype=np.uint8)
        secs = lib().ggref_matrix_time_get(self.h, _p(rows), _p(cols), rows.size, threads,
                                           out.ctypes.data_as(u8p))
        return out, secs


KINDS = {"stat": 0, "sd": 1, "rrr": 2}


class RefBitVector:
    def __init__(self, handle, kind):
        if not handle:
            raise ValueError("reference bit_vector load()/build failed")
        self.h = C.c_void_p(handle)
        self.kind = kind

    @classmethod
    def build(cls, kind, words, n_bits):
        words = np.ascontiguousarray(words, dtype=np.uint64)
        if words.size == 0:
            words = np.zeros(1, dtype=np.uint64)
        return cls(lib().ggref_bv_build(KINDS[kind], _p(words), n_bits), kind)

    @classmethod
    def load(cls, kind, data):
        return cls(lib().ggref_bv_load(KINDS[kind], data, len(data)), kind)

    def serialize(self):
        n = C.c_uint64()
        ptr = lib().ggref_bv_serialize(self.h, C.byref(n))
        return _take_bytes(ptr, n.value)

    def __del__(self):
        if getattr(self, "h", None):
            lib().ggref_bv_free(self.h)
            self.h = None

    size = property(lambda s: lib().ggref_bv_size(s.h))
    num_set_bits = property(lambda s: lib().ggref_bv_num_set_bits(s.h))
    is_inverted = property(lambda s: bool(lib().ggref_bv_is_inverted(s.h)))

    def _q(self, op, q, threads):
        q = np.ascontiguousarray(q, dtype=np.uint64)
        out = np.empty(q.size, dtype=np.uint64)
        secs = lib().ggref_bv_query(self.h, op, _p(q), q.size, _p(out), threads)
        return out, secs

    def rank1(self, q, threads=1):
        return self._q(0, q, threads)[0]

    def select1(self, q, threads=1):
        return self._q(1, q, threads)[0]

    def get(self, q, threads=1):
        return self._q(2, q, threads)[0]

    def _aux(self, op, q, width=64):
        q = np.ascontiguousarray(q, dtype=np.uint64)
        out = np.empty(q.size, dtype=np.uint64)
        f = lib().ggref_bv_query_aux
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p]
        if f(self.h, op, _p(q), q.size, width, _p(out)) != 0:
            raise ValueError("the reference does not offer this operation on %s" % self.kind)
        return out

    def rank0(self, q):
        return self._aux(3, q)

    def select0(self, q):
        return self._aux(4, q)

    def get_int(self, q, width):
        return self._aux(5, q, width)

    def call_ones(self):
        out = np.empty(max(int(self.num_set_bits), 1), dtype=np.uint64)
        f = lib().ggref_bv_call_ones
        f.restype = C.c_uint64
        f.argtypes = [C.c_void_p, C.c_void_p]
        k = f(self.h, _p(out))
        return out[:k]

    def timed(self, op, q, threads):
        return self._q({"rank1": 0, "select1": 1, "get": 2}[op], q, threads)


def gen_norepl(n, m, d, seed=42):
    out = np.zeros((m, (n + 63) // 64), dtype=np.uint64)
    lib().ggref_gen_norepl(n, m, d, seed, _p(out))
    return out


def gen_uniform_rows(n, m, d, n_unique, seed=42):
    out = np.zeros((m, (n + 63) // 64), dtype=np.uint64)
    lib().ggref_gen_uniform_rows(n, m, d, n_unique, seed, _p(out))
    return out


def gen_random_ints(count, begin, end, seed=42):
    out = np.empty(count, dtype=np.uint64)
    lib().ggref_gen_random_ints(count, begin, end, seed, _p(out))
    return out


def gen_implicit(n, m, threshold, seed=42):
    """Portable implicit matrix, column-major words."""
    out = np.zeros((m, (n + 63) // 64), dtype=np.uint64)
    lib().ggref_gen_implicit(n, m, threshold, seed, _p(out))
    return out


def max_threads():
    return lib().ggref_max_threads()
