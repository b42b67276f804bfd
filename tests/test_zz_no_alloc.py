"""GPU: ggq_get_column_noalloc — the get_column variant that allocates nothing on the device (include/ggq.h; no
counterpart in the reference).  Its scratch is a fixed device array inside
the library, so it is checked against the existing entry point (ggq_get_column_into) — bit-exact, the bar of every
integer path here — on calls made back to back (the array keeps the previous call's contents), and its size limit must
be a clear error.  Collected last on purpose (file name)."""
import glob
import os

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu
GOLD = os.path.join(ROOT, "tests", "golden")


def both(gm, col, cap):
    """(rows by the existing entry point, rows by the no-alloc entry point)"""
    a = np.full(cap, 0xA5A5A5A5A5A5A5A5, np.uint64)
    b = np.full(cap, 0x5A5A5A5A5A5A5A5A, np.uint64)
    ka = gm.get_column_into(col, a)
    kb = gm.get_column_noalloc(col, b)
    assert ka == kb == gm.get_column_count(col)
    return a[:ka], b[:kb]


@pytest.mark.parametrize("wide", [0, 1], ids=["default_layout", "capacity_layout"])
def test_no_alloc_column_equals_get_column_on_consecutive_calls(ggq, wide):
    """A synthesized tree (C4 column geometry, 4e5 rows): a dense-ish column, then sparser ones, then the first again —
    two and more calls in a row on the shared scratch, each identical to ggq_get_column_into."""
    ggq.set_default_option("wide_slots", wide)
    try:
        gm = ggq.GpuMatrix.synthesize_brwt(400_000, 2652, 0.003, seed=42, arity=4)
    finally:
        ggq.set_default_option("wide_slots", 0)
    try:
        assert gm.slot_bits == (224 if wide else 96)
        cap = 4096
        first = None
        for col in (0, 1, 2651, 1326, 0, 0, 777):
            a, b = both(gm, col, cap)
            assert a.size > 0 and np.array_equal(a, b), col
            assert np.all(np.diff(b.astype(np.int64)) > 0)   # ascending row ids (BinaryMatrix::get_column)
            if col == 0:
                if first is None:
                    first = b.copy()
                assert np.array_equal(b, first)              # the same answer after other columns went through the array
    finally:
        gm.close()


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "brwt_*.npz"))))
def test_no_alloc_column_on_the_golden_trees(ggq, path):
    """Every column of every golden tree, back to back, against the reference's own answers (tests/golden) and the
    existing entry point; empty columns and the empty matrix included."""
    g = np.load(path)
    gm = ggq.GpuMatrix.load_brwt(g["stream"].tobytes())
    try:
        n, m = int(g["stats"][0]), int(g["stats"][1])
        coff, crows = g["col_offsets"], g["col_rows"]
        for c in range(m):
            a, b = both(gm, c, max(n, 1))
            assert np.array_equal(a, b) and np.array_equal(b, crows[coff[c]:coff[c + 1]]), c
        with pytest.raises(ggq.GgqError):
            gm.get_column_noalloc(m, np.zeros(max(n, 1), np.uint64))          # column out of range
        if m and coff[1] - coff[0] > 0:
            with pytest.raises(ggq.GgqError):
                gm.get_column_noalloc(0, np.zeros(0, np.uint64))              # capacity too small
    finally:
        gm.close()


def test_no_alloc_column_rejects_sizes_beyond_its_scratch(ggq):
    """A column with more set rows than the fixed scratch holds is a clear GGQ_ERR_LIMIT (-5), never a partial answer;
    a column of the same matrix that fits is still served, and the existing entry point serves both."""
    lim = ggq.NOALLOC_MAX_COLUMN_ROWS
    n = 3 * lim
    gm = ggq.GpuMatrix.synthesize_brwt(n, 2, 0.5, seed=7, arity=2)
    try:
        cnt = gm.get_column_count(0)
        assert cnt > lim
        out = np.zeros(cnt, np.uint64)
        with pytest.raises(ggq.GgqError) as e:
            gm.get_column_noalloc(0, out)
        assert e.value.code == -5 and "set rows" in str(e.value) and str(lim) in str(e.value)
        assert gm.get_column_into(0, out) == cnt
    finally:
        gm.close()
    gs = ggq.GpuMatrix.synthesize_brwt(n, 3, 0.3, seed=7, arity=2)   # ~0.9 * 2^20 set rows per column: just fits
    try:
        cnt = gs.get_column_count(1)
        assert lim // 2 < cnt <= lim
        a, b = both(gs, 1, lim)
        assert np.array_equal(a, b)
        a, b = both(gs, 2, lim)
        assert np.array_equal(a, b)
    finally:
        gs.close()


def test_no_alloc_column_is_for_brwt_matrices(ggq):
    g = np.load(os.path.join(GOLD, "colmajor_200x30.npz"))
    cm = ggq.GpuMatrix.load_colmajor(g["stream"].tobytes())
    try:
        with pytest.raises(ggq.GgqError):
            cm.get_column_noalloc(0, np.zeros(200, np.uint64))
    finally:
        cm.close()
