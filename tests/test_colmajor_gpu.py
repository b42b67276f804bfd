"""GPU parity of ColMajorCompressed (K4) through the C ABI vs the unmodified reference
(annotation/column_compressed/column_major.cpp:27-88), on the reference's own serialized stream.
Shapes as in tests/test_column_major.cpp:36-246 (empty, one column, all-zero / all-one, mixed)."""
import numpy as np
import pytest

from conftest import random_dense

pytestmark = pytest.mark.gpu


def check(ggq, ref, dense):
    n, m = dense.shape
    cw = ref.pack_columns(dense) if m else np.zeros((0, 1), dtype=np.uint64)
    rm = ref.RefMatrix.build_colmajor(cw, n)
    data = rm.serialize()
    gm = ggq.GpuMatrix.load_colmajor(data)
    try:
        assert gm.consumed == len(data)
        assert (gm.num_rows, gm.num_columns, gm.num_relations) == (rm.num_rows, rm.num_columns, rm.num_relations)
        rows = np.arange(n, dtype=np.uint64)
        off, cols = gm.get_rows_csr(rows)
        eoff, ecols, _ = rm.get_rows(rows)
        assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
        dup = np.concatenate([rows, rows[::3]])
        assert np.array_equal(gm.count_labels(dup), (dense.sum(0) + dense[::3].sum(0)).astype(np.uint64))
        # the annotator's thresholds on top of the histogram (annotate_static.cpp:71-94, annotate.cpp:57-83), also
        # batched across queries — same code path as for BRWT, fed by the column-major row kernel
        cnt = (dense.sum(0) + dense[::3].sum(0)).astype(np.uint64) if m else np.zeros(0, np.uint64)
        for ratio in (0.0, 0.3):
            need = 1 if ratio == 0 else int(np.ceil(dup.size * ratio))
            keep = np.nonzero((cnt != 0) & (cnt >= need))[0]
            lc, lv = gm.get_labels(dup, ratio)
            assert np.array_equal(lc, keep.astype(np.uint64)) and np.array_equal(lv, cnt[keep])
        if n and m:
            qoff = np.array([0, n // 2, n // 2, dup.size], dtype=np.uint64)
            loff, bc, bv = gm.get_labels_batch(dup, qoff, 0.25)
            for q in range(3):
                ec, ev = gm.get_labels(dup[int(qoff[q]):int(qoff[q + 1])], 0.25)
                assert np.array_equal(bc[int(loff[q]):int(loff[q + 1])], ec) and np.array_equal(bv[int(loff[q]):int(loff[q + 1])], ev)
            tc, tv = gm.get_top_labels(dup, 2)
            nz = np.nonzero(cnt)[0]
            order = nz[np.argsort(-cnt[nz].astype(np.int64), kind="stable")][:2]
            assert np.array_equal(tc, order.astype(np.uint64)) and np.array_equal(tv, cnt[order])
        buf = np.empty(max(n, 1), np.uint64)
        for c in range(m):
            assert np.array_equal(gm.get_column(c), rm.get_column(c)), c
            k = gm.get_column_into(c, buf)
            assert k == gm.get_column_count(c) and np.array_equal(buf[:k], rm.get_column(c))
            assert np.array_equal(gm.get_column(c), np.nonzero(dense[:, c])[0].astype(np.uint64))
        if n * m:
            rr, cc = np.meshgrid(rows, np.arange(m, dtype=np.uint64), indexing="ij")
            assert np.array_equal(gm.get_cells(rr.ravel(), cc.ravel()).reshape(n, m), dense)
            assert np.array_equal(gm.has_column(rows, m - 1), dense[:, m - 1])
            sl = gm.slice_rows(rows[:5])
            exp = []
            for t in range(min(5, n)):
                exp.extend(np.nonzero(dense[t])[0].tolist())
                exp.append(2**64 - 1)
            assert sl.tolist() == exp
    finally:
        gm.close()


def test_edge_shapes(ggq, ref):
    check(ggq, ref, np.zeros((0, 0), np.uint8))
    check(ggq, ref, np.zeros((10, 1), np.uint8))
    check(ggq, ref, np.ones((10, 1), np.uint8))
    check(ggq, ref, np.zeros((40, 5), np.uint8))
    check(ggq, ref, np.ones((40, 5), np.uint8))
    i, j = np.meshgrid(np.arange(19), np.arange(19), indexing="ij")
    check(ggq, ref, ((i + 2 * j) % 2).astype(np.uint8))


@pytest.mark.parametrize("n,m,d", [(1000, 33, 0.05), (1000, 64, 0.6), (777, 100, 0.3), (5000, 7, 0.95),
                                   (300, 200, 0.01), (64, 65, 0.5)])
def test_random(ggq, ref, n, m, d):
    dense = random_dense(n, m, d, seed=n + m)
    dense[:, 0] = 0          # one empty column
    dense[:, m - 1] = 1      # one full (inverted) column
    check(ggq, ref, dense)


def test_bad_stream(ggq, ref):
    dense = random_dense(100, 4, 0.3, seed=1)
    data = ref.RefMatrix.build_colmajor(ref.pack_columns(dense), 100).serialize()
    for bad in (b"", data[:20], data[:-1]):
        with pytest.raises(ggq.GgqError):
            ggq.GpuMatrix.load_colmajor(bad)
