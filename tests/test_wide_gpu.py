"""GPU parity of the CAPACITY LAYOUT (ggq_set_default_option("wide_slots", 1): 32-byte slots of 224
positions instead of 16-byte slots of 96).  The layout changes where the bits live, never an answer: every call must
return what the reference returns (golden fixtures, reference builder) and what the default layout returns on the same
matrix — rows in DFS order, columns, cells, label counts, the serialized stream, the image file."""
import glob
import os

import numpy as np
import pytest

from conftest import ROOT, random_dense

pytestmark = pytest.mark.gpu
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture
def wide(ggq):
    """Matrices created inside the test use the capacity layout; the default is restored afterwards."""
    ggq.set_default_option("wide_slots", 1)
    yield ggq
    ggq.set_default_option("wide_slots", 0)


def both_layouts(ggq, make):
    ggq.set_default_option("wide_slots", 0)
    a = make()
    ggq.set_default_option("wide_slots", 1)
    try:
        b = make()
    finally:
        ggq.set_default_option("wide_slots", 0)
    assert a.slot_bits == 96 and b.slot_bits == 224
    return a, b


def assert_same_answers(ggq, a, b, rows, n_col_probes=9):
    assert (a.num_rows, a.num_columns, a.num_relations, a.num_nodes, a.total_column_size) == \
           (b.num_rows, b.num_columns, b.num_relations, b.num_nodes, b.total_column_size)
    eoff, ecols = a.get_rows_csr(rows)
    off, cols = b.get_rows_csr(rows)
    assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
    m = a.num_columns
    for c in range(0, m, max(1, m // n_col_probes)):
        assert np.array_equal(a.get_column(c), b.get_column(c)), c
    if m and rows.size:
        rng = np.random.default_rng(3)
        cc = rng.integers(0, m, rows.size).astype(np.uint64)
        assert np.array_equal(a.get_cells(rows, cc), b.get_cells(rows, cc))
        assert np.array_equal(a.count_labels(rows), b.count_labels(rows))
        ca, na = a.get_labels(rows[:500], 0.1)
        cb, nb = b.get_labels(rows[:500], 0.1)
        assert np.array_equal(ca, cb) and np.array_equal(na, nb)


def test_unknown_default_option_is_an_error(ggq):
    with pytest.raises(ggq.GgqError):
        ggq.set_default_option("no_such_option", 1)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "brwt_*.npz"))))
def test_golden_fixtures_in_the_capacity_layout(wide, path, tmp_path):
    """The reference's own answers (tests/golden) from a matrix loaded into the capacity layout; serialize() writes the
    reference's stream back byte for byte; an image file keeps the layout it was written with."""
    ggq = wide
    g = np.load(path)
    data = g["stream"].tobytes()
    gm = ggq.GpuMatrix.load_brwt(data)
    try:
        assert gm.slot_bits == 224
        st = [int(x) for x in g["stats"]]
        assert (gm.num_rows, gm.num_columns, gm.num_relations, gm.num_nodes, gm.total_column_size) == tuple(st[:5])
        n, m = st[0], st[1]
        off, cols = gm.get_rows_csr(np.arange(n, dtype=np.uint64))
        assert np.array_equal(off, g["row_offsets"]) and np.array_equal(cols, g["row_cols"])
        coff, crows = g["col_offsets"], g["col_rows"]
        for c in range(m):
            assert np.array_equal(gm.get_column(c), crows[coff[c]:coff[c + 1]])
        if n * m:
            shape = tuple(int(x) for x in g["shape"])
            dense = np.unpackbits(g["dense"])[: n * m].reshape(shape)
            rr, cc = np.meshgrid(np.arange(n, dtype=np.uint64), np.arange(m, dtype=np.uint64), indexing="ij")
            assert np.array_equal(gm.get_cells(rr.ravel(), cc.ravel()).reshape(shape), dense)
        assert gm.serialize() == data
        img = str(tmp_path / "wide.ggqimg")
        gm.save_image(img)
    finally:
        gm.close()
    ggq.set_default_option("wide_slots", 0)   # the file decides, not the default
    gi = ggq.GpuMatrix.load_image(img)
    try:
        assert gi.slot_bits == 224
        off, cols = gi.get_rows_csr(np.arange(n, dtype=np.uint64))
        assert np.array_equal(off, g["row_offsets"]) and np.array_equal(cols, g["row_cols"])
        for c in range(0, m, max(1, m // 5)):
            assert np.array_equal(gi.get_column(c), crows[coff[c]:coff[c + 1]])
        assert gi.serialize() == data
    finally:
        gi.close()


@pytest.mark.parametrize("n_shards", [2, 3])
@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "brwt_*.npz"))))
def test_golden_column_shards_in_the_capacity_layout(wide, path, n_shards):
    """Column shards of a capacity-layout matrix hold exactly the reference's ids of their own columns."""
    ggq = wide
    g = np.load(path)
    data = g["stream"].tobytes()
    n = int(g["stats"][0])
    eoff, ecols = g["row_offsets"], g["row_cols"]
    rows = np.arange(n, dtype=np.uint64)
    got = [[] for _ in range(n)]
    for s in range(n_shards):
        sm = ggq.GpuMatrix.load_brwt_shard(data, s, n_shards)
        try:
            assert sm.slot_bits == 224
            off, cols = sm.get_rows_csr(rows)
            _, _, owned = sm.shard_info()
            for t in range(n):
                part = cols[off[t]:off[t + 1]]
                assert np.all(owned[part.astype(np.int64)])
                got[t].append(part)
        finally:
            sm.close()
    for t in range(n):
        mine = np.concatenate(got[t]) if got[t] else np.zeros(0, np.uint64)
        assert np.array_equal(np.sort(mine), np.sort(ecols[eoff[t]:eoff[t + 1]])), t


@pytest.mark.parametrize("n,m,d,arity", [
    (1, 1, 0.5, 4), (97, 2, 0.3, 2), (223, 3, 0.9, 4), (224, 5, 0.9, 4), (225, 5, 0.9, 2), (5000, 37, 0.05, 4),
    (4097, 65, 0.3, 3), (9000, 50, 0.5, 7), (20000, 17, 0.0, 4), (1000, 9, 1.0, 4), (60_000, 317, 0.038, 4),
    (200_000, 2652, 0.003, 4),
])
def test_synthesized_trees_agree_between_layouts(ggq, n, m, d, arity):
    """K6 writes the capacity layout directly; all rows (small) or a random batch with duplicates (large)."""
    a, b = both_layouts(ggq, lambda: ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=42, arity=arity))
    try:
        rows = np.arange(n, dtype=np.uint64) if n <= 20000 else np.random.default_rng(1).integers(0, n, 50_000).astype(np.uint64)
        assert_same_answers(ggq, a, b, rows)
        if n >= 60_000:
            assert b.device_bytes < a.device_bytes   # 8/7 instead of 4/3 of the plain bits (+ tables)
    finally:
        a.close(); b.close()


def test_capacity_layout_against_direct_evaluation(wide, orc):
    """C5 column geometry (3173 columns, 3.8 %, arity 4) at 1e5 rows: rows equal the direct evaluation of the implicit
    matrix by the C oracle; leaf codes on and off."""
    ggq = wide
    n, m, d = 100_000, 3173, 0.038
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=42, arity=4)
    try:
        assert gm.slot_bits == 224
        thr = ggq.density_threshold(d)
        rows = np.random.default_rng(11).integers(0, n, 1500).astype(np.uint64)
        for codes in (1, 0):
            gm.set_option("leaf_codes", codes)
            off, cols = gm.get_rows_csr(rows)
            for t in range(rows.size):
                assert np.array_equal(cols[off[t]:off[t + 1]], orc.synth_row(m, thr, 42, int(rows[t]))), (t, codes)
    finally:
        gm.close()


def test_leaf_codes_in_the_capacity_layout(ggq):
    """A sparse tree that takes leaf codes (C4 column geometry): coded descent == plain descent == default layout."""
    n, m, d = 80_000, 2652, 0.003
    a, b = both_layouts(ggq, lambda: ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=8, arity=4))
    try:
        assert b.leaf_codes["nodes"] > 0 and b.leaf_codes["in_use"]
        rows = np.random.default_rng(2).integers(0, n, 40_000).astype(np.uint64)
        eoff, ecols = a.get_rows_csr(rows)
        for codes in (1, 0):
            b.set_option("leaf_codes", codes)
            off, cols = b.get_rows_csr(rows)
            assert np.array_equal(off, eoff) and np.array_equal(cols, ecols), codes
    finally:
        a.close(); b.close()


@pytest.mark.parametrize("env", [dict(GGQ_STACK_CAP="16"), dict(GGQ_STACK_UNDERESTIMATE="12"), dict(GGQ_WIDE_REC="1")])
def test_kernel_instantiations_in_the_capacity_layout(ggq, monkeypatch, env):
    """The spill instantiation, the safe-mode rerun and the wide group records of K1, in the capacity layout."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    n, m, d = 30_000, 300, 0.05
    a, b = both_layouts(ggq, lambda: ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=5, arity=4))
    try:
        rows = np.random.default_rng(4).integers(0, n, 20_000).astype(np.uint64)
        assert_same_answers(ggq, a, b, rows, n_col_probes=3)
    finally:
        a.close(); b.close()


@pytest.mark.parametrize("kw", [dict(arity=2), dict(arity=5), dict(greedy=True, arity=2, relax=7)])
def test_device_builder_in_the_capacity_layout(wide, ref, kw):
    """K7 (greedy partitioner + relax) builds the reference builder's tree, and the capacity-layout image serializes to
    the reference builder's bytes; relaxed trees have nodes of arity > 4 (continuation groups)."""
    ggq = wide
    n, m = 6000, 41
    dense = random_dense(n, m, 0.07, 19)
    dense[:, 5] = dense[:, 6]   # similar columns: the greedy partitioner pairs them
    cw = ref.pack_columns(dense)
    rm = ref.RefMatrix.build_brwt(cw, n, threads=4, **kw)
    stream = rm.serialize()
    gm = ggq.GpuMatrix.build_brwt(cw, n, greedy=kw.get("greedy", False), arity=kw.get("arity", 2), relax=kw.get("relax", 0))
    try:
        assert gm.slot_bits == 224
        rows = np.arange(n, dtype=np.uint64)
        eoff, ecols, _ = rm.get_rows(rows, threads=4)
        off, cols = gm.get_rows_csr(rows)
        assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
        assert gm.serialize() == stream
        for c in range(0, m, 5):
            assert np.array_equal(gm.get_column(c), rm.get_column(c)), c
    finally:
        gm.close()


def test_reference_stream_with_long_vectors(wide, ref):
    """A reference-built Multi-BRWT at 3e5 rows (vectors of thousands of blocks, select samples in use) loaded into the
    capacity layout: all rows, every column, pipelined host call."""
    ggq = wide
    n, m = 300_000, 24
    dense = random_dense(n, m, 0.02, 23)
    rm = ref.RefMatrix.build_brwt(ref.pack_columns(dense), n, arity=3, threads=4)
    gm = ggq.GpuMatrix.load_brwt(rm.serialize())
    try:
        assert gm.slot_bits == 224
        rows = np.arange(n, dtype=np.uint64)
        eoff, ecols, _ = rm.get_rows(rows, threads=4)
        off, cols = gm.get_rows_csr(rows)
        assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
        for c in range(m):
            assert np.array_equal(gm.get_column(c), rm.get_column(c)), c
        counts = np.zeros(n, np.uint8)
        ids = np.zeros(int(eoff[-1]), np.uint16)
        total = gm.get_rows_compact(rows.astype(np.uint32), counts, ids)
        assert total == eoff[-1] and np.array_equal(counts, np.diff(eoff).astype(np.uint8))
        assert np.array_equal(ids.astype(np.uint64), ecols)
    finally:
        gm.close()
