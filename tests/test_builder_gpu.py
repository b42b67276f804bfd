"""GPU parity of the device builder (ggq_brwt_build_columns: BRWTBottomUpBuilder::build with the basic or the greedy
partitioner + BRWTOptimizer::relax) against the UNMODIFIED reference builder
(annotation/hierarchical_annotation/BRWT_builders.cpp, partitionings.cpp behind oracle/_ref/libggref.so) on identical
columns: the SAME TREE must come out — every node's parent, arity, column count, first column, vector length and
popcount, in level order — and therefore the same answers, order included, and the same SERIALIZED BYTES
(ggq_brwt_serialize re-encodes the device image as the reference's RangePartition + rrr_vector<63> stream)."""
import numpy as np
import pytest

from conftest import random_dense

pytestmark = pytest.mark.gpu


def compare(ggq, ref, cw, n, rows=None, **kw):
    rm = ref.RefMatrix.build_brwt(cw, n, threads=8, **kw)
    stream = rm.serialize()
    g_ref = ggq.GpuMatrix.load_brwt(stream)
    g_dev = ggq.GpuMatrix.build_brwt(cw, n, greedy=kw.get("greedy", False), arity=kw.get("arity", 2), relax=kw.get("relax", 0))
    try:
        assert (g_dev.num_rows, g_dev.num_columns, g_dev.num_relations) == (g_ref.num_rows, g_ref.num_columns, g_ref.num_relations)
        assert g_dev.num_nodes == g_ref.num_nodes == rm.num_nodes, kw
        assert g_dev.total_column_size == g_ref.total_column_size == rm.total_column_size, kw
        a, b = g_dev.tree_shape(), g_ref.tree_shape()
        for key in a:
            assert np.array_equal(a[key], b[key]), (key, kw)
        if rows is None:
            rows = np.arange(n, dtype=np.uint64)
        eoff, ecols, _ = rm.get_rows(rows, threads=8)
        off, cols = g_dev.get_rows_csr(rows)
        assert np.array_equal(off, eoff) and np.array_equal(cols, ecols), kw
        # BRWT::serialize from the device image: the loaded matrix writes back the stream it was loaded from, and the
        # DEVICE-BUILT matrix writes byte for byte what the reference builder's matrix serializes to
        assert g_ref.serialize() == stream, kw
        assert g_dev.serialize() == stream, kw
        m = g_dev.num_columns
        for c in range(0, m, max(1, m // 7)):
            assert np.array_equal(g_dev.get_column(c), rm.get_column(c)), (c, kw)
    finally:
        g_ref.close()
        g_dev.close()


def test_empty_single_and_tiny(ggq, ref):
    compare(ggq, ref, np.zeros((0, 0), np.uint64), 0)
    for n in (1, 63, 64, 65, 200):
        for m in (1, 2, 3, 5):
            dense = random_dense(n, m, 0.4, seed=n + m)
            cw = ref.pack_columns(dense)
            compare(ggq, ref, cw, n, arity=2)
            compare(ggq, ref, cw, n, greedy=True, relax=3)


@pytest.mark.parametrize("kw", [dict(arity=2), dict(arity=3), dict(arity=5), dict(arity=2, relax=4), dict(arity=2, relax=100),
                                dict(greedy=True), dict(greedy=True, relax=3), dict(greedy=True, relax=7),
                                dict(greedy=True, relax=100)])
def test_reference_shapes_grid(ggq, ref, kw):
    # tests/test_BRWT.cpp:152-212, tests/test_BRWT_optimizer.cpp:102-163: 1..19 x 1..19, all-zero / all-one / checker
    sizes = [1, 2, 3, 5, 8, 13, 19]
    for n in sizes:
        for m in sizes:
            i, j = np.meshgrid(np.arange(n), np.arange(m), indexing="ij")
            for dense in (np.zeros((n, m), np.uint8), np.ones((n, m), np.uint8), ((i + 2 * j) % 2).astype(np.uint8)):
                compare(ggq, ref, ref.pack_columns(dense), n, **kw)


@pytest.mark.parametrize("n,m,d,kw", [
    (1000, 65, 0.05, dict(greedy=True, relax=7)),
    (2017, 33, 0.7, dict(greedy=True, relax=4)),
    (777, 100, 0.02, dict(greedy=True, relax=7)),
    (5000, 300, 0.01, dict(greedy=True, relax=5)),
    (5000, 300, 0.01, dict(arity=4, relax=16)),
    (3000, 40, 0.5, dict(greedy=True, relax=100)),
])
def test_random_matrices(ggq, ref, n, m, d, kw):
    dense = random_dense(n, m, d, seed=n + 7 * m)
    compare(ggq, ref, ref.pack_columns(dense), n, **kw)


def test_config1_and_config2_generators(ggq, ref):
    """The reference's own experiment inputs (experiments/main.cpp:216-247, seed 42), reduced so that the CPU builder
    takes seconds: norepl 1e5 x 100 (brwt_greedy_relax_7) and uniform_rows 1e5 x 500 (Multi-BRWT = greedy + relax 7),
    where columns correlate and the greedy pairing has real structure to find."""
    n = 100_000
    compare(ggq, ref, ref.gen_norepl(n, 100, 0.01, seed=42), n, greedy=True, relax=7)
    compare(ggq, ref, ref.gen_uniform_rows(n, 500, 0.01, n // 5, seed=42), n, greedy=True, relax=7)


@pytest.mark.parametrize("n", [1_200_000, 2_600_000, 10_400_000])
def test_row_sampling_above_one_million_rows(ggq, ref, n):
    """Above 10^6 rows the similarities are estimated on a sample drawn with std::mt19937(1) through
    utils::sample_indexes (utils.cpp:777-816): keep-everything-then-shuffle (n < 2e6), Bernoulli (n <= 1e7) and
    draw-and-dedupe (n > 1e7) branches must all reproduce the reference's pairing."""
    m = 24
    cw = ref.gen_uniform_rows(n, m, 0.05, n // 4, seed=7)
    rows = np.random.default_rng(1).integers(0, n, 200_000).astype(np.uint64)
    compare(ggq, ref, cw, n, rows=rows, greedy=True, relax=6)


@pytest.mark.parametrize("n,m,d,relax", [(200_000, 300, 0.01, 7), (1_500_000, 60, 0.02, 5), (50_000, 97, 0.2, 0),
                                         (2_300_000, 40, 0.004, 7)])
def test_implicit_generator_variant_builds_the_same_tree(ggq, ref, n, m, d, relax):
    """ggq_brwt_synthesize_greedy (the builder for the benchmark's implicit matrix: sampled rows generated on the
    device, popcounts from K6's count pass over all rows) against the reference builder on the MATERIALISED columns
    of the same generator: same tree, same rows, same serialized bytes — below and above the 10^6-row sampling limit."""
    thr = ggq.density_threshold(d)
    cw = ref.gen_implicit(n, m, thr, 42)
    rm = ref.RefMatrix.build_brwt(cw, n, arity=2, greedy=True, relax=relax, threads=8)
    stream = rm.serialize()
    g_ref = ggq.GpuMatrix.load_brwt(stream)
    g_imp = ggq.GpuMatrix.synthesize_brwt_greedy(n, m, d, seed=42, relax=relax)
    try:
        a, b = g_imp.tree_shape(), g_ref.tree_shape()
        for key in a:
            assert np.array_equal(a[key], b[key]), key
        assert (g_imp.num_nodes, g_imp.total_column_size, g_imp.num_relations) == (rm.num_nodes, rm.total_column_size, rm.num_relations)
        rows = np.random.default_rng(3).integers(0, n, 100_000).astype(np.uint64)
        eoff, ecols, _ = rm.get_rows(rows, threads=8)
        off, cols = g_imp.get_rows_csr(rows)
        assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
        assert g_imp.serialize() == stream
    finally:
        g_ref.close()
        g_imp.close()
