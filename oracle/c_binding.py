"""TEST INFRASTRUCTURE — ctypes binding of oracle/liboracle_c.so (plain-C restatement, brwt_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import this.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle_c.so")
u64p = C.POINTER(C.c_uint64)


def available():
    return os.path.exists(_LIB_PATH)


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(_LIB_PATH)
    vp, u64, sz = C.c_void_p, C.c_uint64, C.c_size_t
    sig = {
        "orc_bv_load": (vp, [C.c_int, C.c_char_p, sz, C.POINTER(sz)]),
        "orc_bv_from_words": (vp, [u64p, u64]),
        "orc_bv_free": (None, [vp]),
        "orc_bv_size": (u64, [vp]),
        "orc_bv_num_set_bits": (u64, [vp]),
        "orc_bv_inverted": (C.c_int, [vp]),
        "orc_bv_rank1": (u64, [vp, u64]),
        "orc_bv_select1": (u64, [vp, u64]),
        "orc_bv_get": (C.c_int, [vp, u64]),
        "orc_brwt_load": (vp, [C.c_char_p, sz, C.POINTER(sz)]),
        "orc_brwt_free": (None, [vp]),
        "orc_brwt_num_rows": (u64, [vp]),
        "orc_brwt_num_columns": (u64, [vp]),
        "orc_brwt_num_relations": (u64, [vp]),
        "orc_brwt_num_nodes": (u64, [vp]),
        "orc_brwt_total_column_size": (u64, [vp]),
        "orc_brwt_get": (C.c_int, [vp, u64, u64]),
        "orc_brwt_get_row": (u64, [vp, u64, u64p, u64]),
        "orc_brwt_get_column": (u64, [vp, u64, u64p, u64]),
        "orc_brwt_row_touches": (u64, [vp, u64]),
        "orc_brwt_scan_rows": (u64, [vp, u64p, u64]),
        "orc_colmajor_load": (vp, [C.c_char_p, sz, C.POINTER(sz)]),
        "orc_colmajor_free": (None, [vp]),
        "orc_colmajor_num_rows": (u64, [vp]),
        "orc_colmajor_num_columns": (u64, [vp]),
        "orc_colmajor_num_relations": (u64, [vp]),
        "orc_colmajor_get": (C.c_int, [vp, u64, u64]),
        "orc_colmajor_get_row": (u64, [vp, u64, u64p, u64]),
        "orc_colmajor_get_column": (u64, [vp, u64, u64p, u64]),
        "orc_mix64": (u64, [u64]),
        "orc_synth_row": (u64, [u64, u64, u64, u64, u64p]),
        "orc_synth_bits": (None, [u64, u64, u64, u64p]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _p(a):
    return a.ctypes.data_as(u64p)


class OrcBits:
    def __init__(self, h, consumed=0):
        if not h:
            raise ValueError("oracle: bit vector load failed")
        self.h = C.c_void_p(h)
        self.consumed = consumed

    @classmethod
    def load(cls, kind, data):
        used = C.c_size_t()
        return cls(lib().orc_bv_load({"stat": 0, "sd": 1, "rrr": 2}[kind], data, len(data), C.byref(used)), used.value)

    @classmethod
    def from_words(cls, words, n_bits):
        words = np.ascontiguousarray(words, dtype=np.uint64)
        if words.size == 0:
            words = np.zeros(1, dtype=np.uint64)
        return cls(lib().orc_bv_from_words(_p(words), n_bits))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_bv_free(self.h)
            self.h = None

    size = property(lambda s: lib().orc_bv_size(s.h))
    num_set_bits = property(lambda s: lib().orc_bv_num_set_bits(s.h))
    is_inverted = property(lambda s: bool(lib().orc_bv_inverted(s.h)))

    def rank1(self, q):
        return np.array([lib().orc_bv_rank1(self.h, int(x)) for x in q], dtype=np.uint64)

    def select1(self, q):
        return np.array([lib().orc_bv_select1(self.h, int(x)) for x in q], dtype=np.uint64)

    def get(self, q):
        return np.array([lib().orc_bv_get(self.h, int(x)) for x in q], dtype=np.uint64)


class _OrcMatrix:
    _pfx = None

    def __init__(self, data):
        used = C.c_size_t()
        h = getattr(lib(), self._pfx + "_load")(data, len(data), C.byref(used))
        if not h:
            raise ValueError("oracle: load failed")
        self.h = C.c_void_p(h)
        self.consumed = used.value

    def __del__(self):
        if getattr(self, "h", None):
            getattr(lib(), self._pfx + "_free")(self.h)
            self.h = None

    def _f(self, name):
        return getattr(lib(), self._pfx + "_" + name)

    num_rows = property(lambda s: s._f("num_rows")(s.h))
    num_columns = property(lambda s: s._f("num_columns")(s.h))
    num_relations = property(lambda s: s._f("num_relations")(s.h))

    def get(self, row, col):
        return bool(self._f("get")(self.h, row, col))

    def get_row(self, row):
        cap = max(1, int(self.num_columns))
        out = np.empty(cap, dtype=np.uint64)
        k = self._f("get_row")(self.h, row, _p(out), cap)
        return out[:k].copy()

    def get_column(self, col):
        cap = max(1, int(self.num_rows))
        out = np.empty(cap, dtype=np.uint64)
        k = self._f("get_column")(self.h, col, _p(out), cap)
        return out[:k].copy()


class OrcBRWT(_OrcMatrix):
    _pfx = "orc_brwt"
    num_nodes = property(lambda s: lib().orc_brwt_num_nodes(s.h))
    total_column_size = property(lambda s: lib().orc_brwt_total_column_size(s.h))

    def row_touches(self, row):
        return lib().orc_brwt_row_touches(self.h, row)

    def scan_rows(self, rows):
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        return lib().orc_brwt_scan_rows(self.h, _p(rows), rows.size)


class OrcColMajor(_OrcMatrix):
    _pfx = "orc_colmajor"


def mix64(x):
    return lib().orc_mix64(x)


def synth_row(n_cols, threshold, seed, row):
    out = np.empty(max(1, n_cols), dtype=np.uint64)
    k = lib().orc_synth_row(n_cols, threshold, seed, row, _p(out))
    return out[:k].copy()


def synth_bits(n_bits, threshold, seed):
    words = np.zeros((n_bits + 63) // 64, dtype=np.uint64)
    if n_bits:
        lib().orc_synth_bits(n_bits, threshold, seed, _p(words))
    return words
