"""GPU parity: BRWT / Multi-BRWT queries through the C ABI (libggq.so) vs the unmodified reference
(oracle/_ref/libggref.so: BRWT::get_row / get / get_column, BRWT.cpp:9-85) on identical serialized
streams.  Bit-exact, order included (get_row order = DFS over children).

Shapes follow the reference's own tests (tests/test_BRWT.cpp:152-212: 1..19 x 1..19 all-zero / all-one /
(i+2j)%2; tests/test_BRWT_optimizer.cpp:102-163: the same after relax) plus the SURVEY §4 grid.
"""
import numpy as np
import pytest

from conftest import random_dense

pytestmark = pytest.mark.gpu


def build_ref(ref, dense, **kw):
    n, m = dense.shape
    cw = ref.pack_columns(dense) if m else np.zeros((0, 0), dtype=np.uint64)
    return ref.RefMatrix.build_brwt(cw, n, **kw)


def check_matrix(ggq, rm, dense=None, rows=None, all_cells=True):
    data = rm.serialize()
    gm = ggq.GpuMatrix.load_brwt(data)
    try:
        assert gm.consumed == len(data)
        n, m = rm.num_rows, rm.num_columns
        assert (gm.num_rows, gm.num_columns, gm.num_relations) == (n, m, rm.num_relations)
        assert gm.num_nodes == rm.num_nodes
        assert gm.total_column_size == rm.total_column_size
        if rows is None:
            rows = np.arange(n, dtype=np.uint64)
        rows = np.asarray(rows, dtype=np.uint64)
        eoff, ecols, _ = rm.get_rows(rows)
        off, cols = gm.get_rows_csr(rows)
        assert np.array_equal(off, eoff)
        assert np.array_equal(cols, ecols)
        # count_labels: histogram over a multiset of rows (annotate_static.cpp:149-162), duplicates counted
        half = rows[:rows.size // 2]
        exp_cnt = np.bincount(ecols.astype(np.int64), minlength=m) + \
            np.bincount(ecols[:int(eoff[half.size])].astype(np.int64), minlength=m)
        assert np.array_equal(gm.count_labels(np.concatenate([rows, half])), exp_cnt.astype(np.uint64)[:m])
        # get_labels(indices, ratio) (annotate_static.cpp:71-94) and get_top_labels (annotate.cpp:57-83)
        both = np.concatenate([rows, half])
        for ratio in (0.0, 0.01, 0.25, 0.5, 1.0):
            need = 1 if ratio == 0 else int(np.ceil(both.size * ratio))
            keep = np.nonzero((exp_cnt[:m] != 0) & (exp_cnt[:m] >= need))[0]
            lc, lv = gm.get_labels(both, ratio)
            assert np.array_equal(lc, keep.astype(np.uint64)) and np.array_equal(lv, exp_cnt[keep].astype(np.uint64))
        for num_top in (0, 1, 3, m, m + 5):
            nzc = np.nonzero(exp_cnt[:m])[0]
            order = nzc[np.argsort(-exp_cnt[nzc].astype(np.int64), kind="stable")][:num_top]
            tc, tv = gm.get_top_labels(both, num_top)
            assert np.array_equal(tc, order.astype(np.uint64)) and np.array_equal(tv, exp_cnt[order].astype(np.uint64))
        # slice_rows = rows separated by UINT64_MAX
        sl = gm.slice_rows(rows)
        exp = []
        for t in range(rows.size):
            exp.extend(ecols[eoff[t]:eoff[t + 1]].tolist())
            exp.append(2**64 - 1)
        assert np.array_equal(sl, np.array(exp, dtype=np.uint64))
        for c in range(m):
            assert np.array_equal(gm.get_column(c), rm.get_column(c)), c
        if all_cells and n * m:
            rr, cc = np.meshgrid(np.arange(n, dtype=np.uint64), np.arange(m, dtype=np.uint64), indexing="ij")
            got = gm.get_cells(rr.ravel(), cc.ravel())
            if dense is not None:
                assert np.array_equal(got.reshape(n, m), dense)
            else:
                e, _ = rm.get_cells(rr.ravel(), cc.ravel())
                assert np.array_equal(got, e)
            for c in range(0, m, max(1, m // 5)):
                assert np.array_equal(gm.has_column(rows, c), got.reshape(n, m)[rows.astype(np.int64), c])
    finally:
        gm.close()


def test_empty_and_single(ggq, ref):
    # tests/test_BRWT.cpp:15-90: empty, one column (root is a leaf), two columns
    check_matrix(ggq, build_ref(ref, np.zeros((0, 0), np.uint8)))
    for n in (1, 5, 63, 64, 97):
        for pat in (0, 1, 2):
            d = np.zeros((n, 1), np.uint8) if pat == 0 else np.ones((n, 1), np.uint8)
            if pat == 2:
                d[::2] = 0
            check_matrix(ggq, build_ref(ref, d), d)
    d = np.array([[1, 0], [0, 1], [1, 1], [0, 0], [1, 0]], np.uint8)
    check_matrix(ggq, build_ref(ref, d), d)


@pytest.mark.parametrize("relax", [0, 100])
def test_small_shape_grid(ggq, ref, relax):
    # shapes 1..19 x 1..19 (subsampled odd sizes + 19), three fill patterns, arity 2
    sizes = [1, 2, 3, 5, 8, 13, 19]
    for n in sizes:
        for m in sizes:
            i, j = np.meshgrid(np.arange(n), np.arange(m), indexing="ij")
            for dense in (np.zeros((n, m), np.uint8), np.ones((n, m), np.uint8), ((i + 2 * j) % 2).astype(np.uint8),
                          ((i + j) % 2).astype(np.uint8)):
                check_matrix(ggq, build_ref(ref, dense, arity=2, relax=relax), dense)


@pytest.mark.parametrize("n,m,d,kw", [
    (1000, 1, 0.3, dict(arity=2)),
    (1000, 2, 0.3, dict(arity=2)),
    (1000, 63, 0.05, dict(arity=2)),
    (1000, 64, 0.05, dict(arity=4)),
    (1000, 65, 0.05, dict(arity=3)),
    (1000, 65, 0.5, dict(arity=5)),
    (777, 100, 0.02, dict(arity=2, greedy=True, relax=7)),
    (2017, 33, 0.7, dict(arity=2, greedy=True, relax=4)),
    (96 * 3 + 1, 20, 1.0, dict(arity=4)),
    (96 * 3, 20, 0.0, dict(arity=4)),
    (500, 300, 0.01, dict(arity=7)),
    (500, 300, 0.01, dict(arity=40)),
])
def test_random_shapes(ggq, ref, n, m, d, kw):
    dense = random_dense(n, m, d, seed=n * 31 + m)
    check_matrix(ggq, build_ref(ref, dense, **kw), dense)


def test_wide_1000_columns(ggq, ref):
    dense = random_dense(1000, 1000, 0.01, seed=5)
    rm = build_ref(ref, dense, arity=2, greedy=True, relax=7)
    check_matrix(ggq, rm, dense, all_cells=True)


def test_duplicate_and_out_of_range_rows(ggq, ref):
    dense = random_dense(300, 40, 0.1, seed=3)
    rm = build_ref(ref, dense, arity=4)
    gm = ggq.GpuMatrix.load_brwt(rm.serialize())
    rows = np.array([5, 5, 299, 0, 5, 17, 17], dtype=np.uint64)
    off, cols = gm.get_rows_csr(rows)
    eoff, ecols, _ = rm.get_rows(rows)
    assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
    # rows >= num_rows are UB in the reference (assert only); the engine documents "empty row"
    off, cols = gm.get_rows_csr(np.array([300, 2**40, 3], dtype=np.uint64))
    assert off[1] == 0 and off[2] == 0
    assert np.array_equal(cols, rm.get_row(3))
    off, cols = gm.get_rows_csr(np.zeros(0, dtype=np.uint64))
    assert off.tolist() == [0] and cols.size == 0
    gm.close()


def test_bad_stream_is_rejected(ggq, ref):
    dense = random_dense(200, 10, 0.2, seed=9)
    data = build_ref(ref, dense, arity=2).serialize()
    for bad in (b"", data[:50], data[: len(data) // 2], b"\xff" * 64):
        with pytest.raises(ggq.GgqError) as e:
            ggq.GpuMatrix.load_brwt(bad)
        assert e.value.code == -1  # GGQ_ERR_FORMAT  <->  BRWT::load() == false (BRWT.cpp:108-110)


def test_config1_norepl_1e6(ggq, ref):
    # BASELINE config 1 at its full size: experiments/main.cpp:216-217 generator (seed 42), 1e6 x 100, brwt_greedy_relax_7
    n, m = 1_000_000, 100
    cw = ref.gen_norepl(n, m, 0.01, seed=42)
    rm = ref.RefMatrix.build_brwt(cw, n, arity=2, greedy=True, relax=7, threads=16)
    gm = ggq.GpuMatrix.load_brwt(rm.serialize())
    rows = np.arange(n, dtype=np.uint64)
    off, cols = gm.get_rows_csr(rows)
    eoff, ecols, _ = rm.get_rows(rows, threads=8)
    assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
    # identity of SURVEY §6: every bit of every node vector is visited exactly once
    for c in (0, 17, 99):
        assert np.array_equal(gm.get_column(c), rm.get_column(c))
    rnd = ref.gen_random_ints(1000, 0, n, seed=42)
    off, cols = gm.get_rows_csr(rnd)
    eoff, ecols, _ = rm.get_rows(rnd)
    assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
    gm.close()


def test_config2_uniform_rows_multi_brwt(ggq, ref):
    """BASELINE config 2 (SURVEY 8d "C2 concrete input") at its full size, 1e6 x 1000:
    generate_random_rows(2e5 unique x 5, shuffled) seed 42 (experiments/main.cpp:232-247), Multi-BRWT = greedy
    partitioner + relax(7) (the reference builder takes about a minute on the box's host threads); batched get_rows
    over all rows and over random ids at batch sizes 1 / 1e2 / 1e4 / all, get_column for 10 random columns and a
    sweep of the others."""
    n, m = 1_000_000, 1000
    cw = ref.gen_uniform_rows(n, m, 0.01, n // 5, seed=42)
    rm = ref.RefMatrix.build_brwt(cw, n, arity=2, greedy=True, relax=7, threads=16)
    gm = ggq.GpuMatrix.load_brwt(rm.serialize())
    try:
        assert (gm.num_rows, gm.num_columns, gm.num_relations) == (n, m, rm.num_relations)
        assert gm.num_nodes == rm.num_nodes and gm.total_column_size == rm.total_column_size
        rows = np.arange(n, dtype=np.uint64)
        eoff, ecols, _ = rm.get_rows(rows, threads=8)
        off, cols = gm.get_rows_csr(rows)
        assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
        rnd = ref.gen_random_ints(n, 0, n, seed=42)
        roff, rcols, _ = rm.get_rows(rnd, threads=8)
        for bs in (1, 100, 10_000, n):
            lim = min(n, bs * 20)   # 20 batches of each size
            for lo in range(0, lim, bs):
                o, c = gm.get_rows_csr(rnd[lo:lo + bs])
                assert np.array_equal(o, roff[lo:lo + bs + 1] - roff[lo])
                assert np.array_equal(c, rcols[int(roff[lo]):int(roff[lo + bs])])
        assert np.array_equal(gm.count_labels(rnd), np.bincount(rcols.astype(np.int64), minlength=m).astype(np.uint64))
        some = ref.gen_random_ints(10, 0, m, seed=42).tolist() + list(range(0, m, 37))
        for c in some:
            assert np.array_equal(gm.get_column(int(c)), rm.get_column(int(c)))
    finally:
        gm.close()


@pytest.mark.parametrize("rb", [4, 8, 16])
@pytest.mark.parametrize("m,relax,seed", [(825, 8, 1), (422, 5, 2), (1500, 3, 3)])
def test_irregular_trees_random_row_densities(ggq, ref, m, relax, seed, rb):
    """Irregular Multi-BRWT shapes (greedy pairing + relax: arities 2..relax, leaves at different depths) queried with
    rows whose densities differ wildly (empty .. nearly full): the shape that can stack higher in the descent kernel
    than the host's replay predicts.  Every row against the unmodified reference, for 4 / 8 / 16 rows per warp tile."""
    n = 1500
    rng = np.random.default_rng(seed)
    row_d = rng.choice([0.0, 0.002, 0.02, 0.2, 0.6, 0.97], size=n)
    col_d = rng.choice([0.2, 1.0, 1.0, 1.5], size=m)
    dense = (rng.random((n, m)) < np.minimum(1.0, row_d[:, None] * col_d[None, :])).astype(np.uint8)
    rm = build_ref(ref, dense, arity=2, greedy=True, relax=relax, threads=8)
    gm = ggq.GpuMatrix.load_brwt(rm.serialize())
    try:
        gm.set_option("rows_per_warp", rb)
        rows = rng.integers(0, n, 20_000).astype(np.uint64)
        eoff, ecols, _ = rm.get_rows(rows, threads=8)
        off, cols = gm.get_rows_csr(rows)
        assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
        assert np.array_equal(gm.count_labels(rows), np.bincount(ecols.astype(np.int64), minlength=m).astype(np.uint64))
    finally:
        gm.close()


def test_leaf_codes_on_reference_built_multi_brwt(ggq, ref):
    """Leaf codes on a tree made by the reference's own builder (greedy pairing + relax: irregular arities, leaves at
    different depths, non-identity column arrangement): rows with codes on and off against the unmodified reference."""
    n, m = 60_000, 600
    cw = ref.gen_norepl(n, m, 0.004, seed=7)
    rm = ref.RefMatrix.build_brwt(cw, n, arity=2, greedy=True, relax=6, threads=8)
    gm = ggq.GpuMatrix.load_brwt(rm.serialize())
    try:
        info = gm.leaf_codes
        assert info["nodes"] > 0 and info["in_use"], info
        rows = ref.gen_random_ints(50_000, 0, n, seed=3)
        eoff, ecols, _ = rm.get_rows(rows, threads=8)
        for on in (1, 0):
            gm.set_option("leaf_codes", on)
            off, cols = gm.get_rows_csr(rows)
            assert np.array_equal(off, eoff) and np.array_equal(cols, ecols), on
            lc, lv = gm.get_labels(rows[:300], 0.0)
            h = np.bincount(ecols[:int(eoff[300])].astype(np.int64), minlength=m)
            assert np.array_equal(lc, np.nonzero(h)[0].astype(np.uint64)) and np.array_equal(lv, h[h != 0].astype(np.uint64))
    finally:
        gm.close()
