"""GPU parity of the device-side BRWT builder (harness for configs C4/C5): the tree
synthesized on the device for the implicit matrix M[i][j] = mix64(seed*K + i*m + j) < T must answer every
query exactly like the reference BRWT that BRWTBottomUpBuilder::build(columns, get_basic_partitioner(arity))
(BRWT_builders.cpp:19-31,122-163) makes from the same, materialised matrix."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def implicit_dense(ggq, n, m, thr, seed):
    i = np.arange(n, dtype=np.uint64)[:, None]
    j = np.arange(m, dtype=np.uint64)[None, :]
    with np.errstate(over="ignore"):
        x = np.uint64((seed * ggq.SEED_MUL) % 2**64) + i * np.uint64(m) + j
    return (ggq.mix64(x) < np.uint64(thr)).astype(np.uint8)


@pytest.mark.parametrize("n,m,d,arity", [
    (1, 1, 0.5, 4), (100, 1, 0.3, 4), (97, 2, 0.3, 2), (5000, 37, 0.05, 4), (4096, 64, 0.003, 4),
    (4097, 65, 0.3, 3), (10000, 100, 0.01, 2), (3000, 266, 0.038, 4), (20000, 17, 0.0, 4), (1000, 9, 1.0, 4),
    (9000, 50, 0.5, 7),
])
def test_synth_matches_reference_builder(ggq, ref, n, m, d, arity):
    thr = ggq.density_threshold(d)
    dense = implicit_dense(ggq, n, m, thr, 42)
    rm = ref.RefMatrix.build_brwt(ref.pack_columns(dense), n, arity=arity)
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=42, arity=arity)
    try:
        assert (gm.num_rows, gm.num_columns) == (n, m)
        assert gm.num_relations == rm.num_relations == int(dense.sum())
        assert gm.num_nodes == rm.num_nodes
        assert gm.total_column_size == rm.total_column_size
        rows = np.arange(n, dtype=np.uint64)
        off, cols = gm.get_rows_csr(rows)
        eoff, ecols, _ = rm.get_rows(rows)
        assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
        for c in range(0, m, max(1, m // 16)):
            assert np.array_equal(gm.get_column(c), rm.get_column(c)), c
        rr = np.repeat(rows, min(m, 8))
        cc = np.tile(np.arange(min(m, 8), dtype=np.uint64) * np.uint64(max(1, m // 8)), n)
        assert np.array_equal(gm.get_cells(rr, cc), dense[rr.astype(np.int64), cc.astype(np.int64)])
    finally:
        gm.close()


def test_synth_kingsford_shaped_sample(ggq, orc):
    """2652 columns, d = 0.3 %, arity 4 (C4 shape) at 2e5 rows: every sampled row equals the direct
    evaluation of M[i][.] by the C oracle (orc_synth_row), in ascending = DFS order for the basic partitioner."""
    n, m, d = 200_000, 2652, 0.003
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=42, arity=4)
    thr = ggq.density_threshold(d)
    rng = np.random.default_rng(7)
    rows = rng.integers(0, n, 4000).astype(np.uint64)
    off, cols = gm.get_rows_csr(rows)
    for t in range(0, rows.size):
        e = orc.synth_row(m, thr, 42, int(rows[t]))
        assert np.array_equal(cols[off[t]:off[t + 1]], e), t
    assert gm.num_nodes == 3538  # SURVEY §6 probe: arity-4 basic tree over 2652 columns
    gm.close()
