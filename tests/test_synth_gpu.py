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


def test_row_query_options_do_not_change_results(ggq, orc):
    """The engine may reorder a batch internally (radix sort on the top bits of the row id) and pick any
    rows-per-warp tile; the CSR must stay in the caller's order and bit-identical.  Duplicates and
    out-of-range ids (empty rows, ggq.h) included."""
    n, m, d = 30_000, 300, 0.02
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=42, arity=3)
    thr = ggq.density_threshold(d)
    rng = np.random.default_rng(11)
    rows = rng.integers(0, n, 20_000).astype(np.uint64)
    rows[::97] = rows[5]                      # duplicates
    rows[::1013] = np.uint64(n + 12345)       # out of range
    rows[7] = np.uint64(2**63)
    base = None
    try:
        for sort_bits in (0, 5, 64):
            for rb in (1, 4, 16):
                gm.set_option("sort_min_rows", 0)
                gm.set_option("sort_bits", sort_bits)
                gm.set_option("rows_per_warp", rb)
                off, cols = gm.get_rows_csr(rows)
                if base is None:
                    base = (off, cols)
                    for t in range(0, rows.size, 7):
                        r = int(rows[t])
                        e = orc.synth_row(m, thr, 42, r) if r < n else np.zeros(0, np.uint64)
                        assert np.array_equal(cols[off[t]:off[t + 1]], e), t
                else:
                    assert np.array_equal(off, base[0]) and np.array_equal(cols, base[1]), (sort_bits, rb)
                sl = gm.slice_rows(rows[:5000])
                assert sl.size == int(off[5000]) + 5000
    finally:
        gm.close()


def test_pipelined_host_api_matches_single_shot(ggq):
    """ggq_get_rows_into cuts batches above 2^20 rows into chunks on two streams; the CSR must equal the
    concatenation of the unpipelined answers (ggq_get_rows), global offsets included."""
    n, m, d = 60_000, 120, 0.03
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=7, arity=4)
    try:
        rng = np.random.default_rng(3)
        rows = rng.integers(0, n, (1 << 21) + 12_345).astype(np.uint64)
        off = np.empty(rows.size + 1, np.uint64)
        cols = np.empty(int(rows.size * m * d * 1.3) + 1024, np.uint64)
        total = gm.get_rows_into(rows, off, cols)
        assert off[0] == 0 and off[-1] == total
        assert np.all(np.diff(off.astype(np.int64)) >= 0)
        for lo in range(0, rows.size, 700_000):
            hi = min(rows.size, lo + 700_000)
            eo, ec = gm.get_rows_csr(rows[lo:hi])
            assert np.array_equal(off[lo:hi + 1] - off[lo], eo)
            assert np.array_equal(cols[int(off[lo]):int(off[hi])], ec)
        # count_labels above 2^22 rows runs the same chunk pipeline (histograms of all chunks add up)
        hist = np.bincount(cols[:total].astype(np.int64), minlength=m).astype(np.uint64)
        assert np.array_equal(gm.count_labels(rows), hist)
        assert np.array_equal(gm.count_labels(np.tile(rows, 3)), hist * np.uint64(3))
        # >= 4 chunks: the ramped chunk plan (quarter / half chunks at both ends)
        rows3 = np.tile(rows, 3)
        off3 = np.empty(rows3.size + 1, np.uint64)
        cols3 = np.empty(3 * total + 16, np.uint32)
        assert gm.get_rows_into(rows3, off3, cols3) == 3 * total
        for rep in range(3):
            lo = rep * rows.size
            assert np.array_equal(off3[lo:lo + rows.size + 1] - off3[lo], off)
            assert np.array_equal(cols3[int(off3[lo]):int(off3[lo + rows.size])], cols[:total].astype(np.uint32))
        # out-of-range ids in a pipelined batch (empty rows, ggq.h): the counting sort of the chunks clamps them
        # into its last bucket
        bad = rows.copy()
        bad[::4099] = np.uint64(n + 7)
        bad[5::8191] = np.uint64(2**64 - 2)
        offb = np.empty_like(off)
        colsb = np.empty(cols.size, np.uint32)
        totb = gm.get_rows_into(bad, offb, colsb)
        lens, lensb = np.diff(off.astype(np.int64)), np.diff(offb.astype(np.int64))
        is_bad = bad >= np.uint64(n)
        assert np.all(lensb[is_bad] == 0) and np.array_equal(lensb[~is_bad], lens[~is_bad])
        keep = np.repeat(~is_bad, lens)
        assert np.array_equal(colsb[:totb], cols[:total][keep].astype(np.uint32))
        # compact u32 ids: same CSR
        cols32 = np.empty(cols.size, np.uint32)
        off32 = np.empty_like(off)
        assert gm.get_rows_into(rows, off32, cols32) == total
        assert np.array_equal(off32, off) and np.array_equal(cols32[:total], cols[:total].astype(np.uint32))
        # capacity too small: error, but the total is still reported
        small = np.empty(10, np.uint64)
        with pytest.raises(ggq.GgqError):
            gm.get_rows_into(rows, off, small)
    finally:
        gm.close()


def test_two_matrices_with_different_shapes_coexist(ggq):
    """Kernel attributes (dynamic shared memory limit) are per function, not per matrix: a wide matrix must
    still answer after a narrow one was created (regression)."""
    wide = ggq.GpuMatrix.synthesize_brwt(5000, 2500, 0.01, seed=1, arity=4)
    rows = np.arange(0, 5000, 3, dtype=np.uint64)
    a = wide.get_rows_csr(rows)
    narrow = ggq.GpuMatrix.synthesize_brwt(5000, 8, 0.2, seed=1, arity=2)
    narrow.get_rows_csr(rows)
    b = wide.get_rows_csr(rows)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    wide.close(); narrow.close()


def _expected_labels(orc, m, thr, seed, rows, n, ratio):
    """StaticBinRelAnnotator::get_labels(indices, ratio) (annotate_static.cpp:71-94) restated on direct evaluation of
    the implicit matrix (oracle): histogram over the multiset of rows, min = ratio == 0 ? 1 : ceil(|rows| * ratio)."""
    h = np.zeros(m, np.int64)
    for r in rows:
        if r < n:
            np.add.at(h, orc.synth_row(m, thr, seed, int(r)).astype(np.int64), 1)
    need = 1 if ratio == 0 else int(np.ceil(len(rows) * ratio))
    keep = np.nonzero((h != 0) & (h >= need))[0]
    return keep.astype(np.uint64), h[keep].astype(np.uint64)


@pytest.mark.parametrize("sizes", [[0, 1, 2, 300, 17, 0, 250, 299], "many", "large"])
def test_labels_batch_against_oracle_histograms(ggq, orc, sizes):
    """ggq_get_labels_batch / _ratios (cross-query batching, main.cpp:462-504): every query's kept labels and counts
    against histograms of oracle-evaluated rows and the reference's threshold rule, with one ratio for all queries
    and with a ratio per query (AnnotatedDBG::get_labels rescales it per sequence, annotated_dbg.cpp:104-113).
    Covers the direct small-batch form (<= 2^14 indices), the general form, empty queries and duplicates."""
    n, m, d, seed = 40_000, 150, 0.04, 11
    thr = ggq.density_threshold(d)
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=seed, arity=3)
    try:
        rng = np.random.default_rng(5)
        if sizes == "many":
            sizes = np.concatenate([[0, 1, 2], rng.integers(1, 200, 100), [0, 900]])       # ~11 K indices: direct form
        elif sizes == "large":
            sizes = np.concatenate([[0, 3], rng.integers(100, 400, 90), [0, 6000]])         # ~28 K indices: general form
        sizes = np.asarray(sizes)
        qoff = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
        rows = rng.integers(0, n, int(qoff[-1])).astype(np.uint64)
        if rows.size > 20:
            rows[10:20] = rows[10]   # duplicates are counted
            rows[5] = np.uint64(n + 7)   # out of range: an empty row
        per_query = rng.choice([0.0, 0.01, 0.05, 0.3, 1.0], size=sizes.size)
        for ratio in (0.0, 0.05, 0.3, 1.0, per_query):
            loff, cols, cnt = gm.get_labels_batch(rows, qoff, ratio)
            assert loff[0] == 0 and loff[-1] == cols.size == cnt.size
            for q in range(sizes.size):
                rq = float(ratio[q]) if np.ndim(ratio) else ratio
                ec, ev = _expected_labels(orc, m, thr, seed, rows[int(qoff[q]):int(qoff[q + 1])], n, rq)
                a, b = int(loff[q]), int(loff[q + 1])
                assert np.array_equal(cols[a:b], ec) and np.array_equal(cnt[a:b], ev), (rq, q)
        # the single-query call agrees as well
        for q in (3, sizes.size - 1):
            sub = rows[int(qoff[q]):int(qoff[q + 1])]
            lc, lv = gm.get_labels(sub, 0.05)
            ec, ev = _expected_labels(orc, m, thr, seed, sub, n, 0.05)
            assert np.array_equal(lc, ec) and np.array_equal(lv, ev)
        # capacity too small: error, offsets complete
        if qoff[-1] > 100:
            with pytest.raises(ggq.GgqError):
                gm.get_labels_batch(rows, qoff, 0.0, capacity=3)
        # no queries
        loff, cols, cnt = gm.get_labels_batch(np.zeros(0, np.uint64), np.zeros(1, np.uint64), 0.5)
        assert loff.tolist() == [0] and cols.size == 0
    finally:
        gm.close()


def test_concurrent_const_calls_from_many_threads(ggq):
    """The reference's callers run const queries concurrently from a thread pool on ONE matrix object
    (main.cpp:462-491); the engine answers every calling thread on its own stream + scratch (workspace pool).
    ctypes releases the GIL during the calls, so these really overlap."""
    import threading
    n, m, d = 80_000, 200, 0.03
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=21, arity=4)
    try:
        rng = np.random.default_rng(9)
        batches = [rng.integers(0, n, int(sz)).astype(np.uint64) for sz in rng.integers(1, 60_000, 24)]
        expect = [gm.get_rows_csr(b) for b in batches]
        expect_cnt = [gm.count_labels(b) for b in batches]
        expect_lab = [gm.get_labels(b, 0.02) for b in batches]
        errors = []

        def worker(tid):
            try:
                for rep in range(3):
                    for i in range(tid, len(batches), 4):
                        off, cols = gm.get_rows_csr(batches[i])
                        assert np.array_equal(off, expect[i][0]) and np.array_equal(cols, expect[i][1])
                        assert np.array_equal(gm.count_labels(batches[i]), expect_cnt[i])
                        lc, lv = gm.get_labels(batches[i], 0.02)
                        assert np.array_equal(lc, expect_lab[i][0]) and np.array_equal(lv, expect_lab[i][1])
                        col = int(batches[i][0]) % m
                        hc = gm.has_column(batches[i][:1000], col)
                        ex = np.array([col in cols[off[t]:off[t + 1]] for t in range(min(1000, batches[i].size))])
                        assert np.array_equal(hc.astype(bool), ex)
            except Exception as e:   # noqa: BLE001
                errors.append((tid, repr(e)))

        threads = [threading.Thread(target=worker, args=(t,)) for t in range(8)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        assert not errors, errors
    finally:
        gm.close()


def test_stack_spill_instantiation_gives_the_same_rows(ggq, orc, monkeypatch):
    """Trees whose worst-case descent stack exceeds the shared-memory budget use the SPILL instantiation of the
    descent kernel (overflow area in global memory).  GGQ_STACK_CAP shrinks the shared part so that an ordinary
    tree takes that path; dense rows make the stack actually overflow into the global part."""
    n, m, d = 20_000, 400, 0.35
    rows = np.random.default_rng(4).integers(0, n, 30_000).astype(np.uint64)
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=5, arity=4)
    base = gm.get_rows_csr(rows)
    gm.close()
    thr = ggq.density_threshold(d)
    for t in range(0, rows.size, 997):
        assert np.array_equal(base[1][base[0][t]:base[0][t + 1]], orc.synth_row(m, thr, 5, int(rows[t])))
    monkeypatch.setenv("GGQ_STACK_CAP", "16")
    gs = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=5, arity=4)
    try:
        off, cols = gs.get_rows_csr(rows)
        assert np.array_equal(off, base[0]) and np.array_equal(cols, base[1])
        assert np.array_equal(gs.count_labels(rows), np.bincount(base[1].astype(np.int64), minlength=m).astype(np.uint64))
    finally:
        gs.close()


@pytest.mark.parametrize("n,m,d,arity", [(20_000, 400, 0.02, 4), (50_000, 2652, 0.003, 4), (9_000, 36, 0.4, 2),
                                         (3_000, 1, 0.5, 4)])
def test_flat_emit_gives_the_same_rows(ggq, n, m, d, arity, monkeypatch):
    """EMIT_FLAT instantiation of the descent kernel (flat 128-bit scan over the tile's bitmaps, no emit queue) must
    return exactly the rows of the default emit; where the bitmap words per tile are not a multiple of 4 the engine
    keeps the default emit, which this test then covers trivially."""
    rows = np.random.default_rng(9).integers(0, n, 40_000).astype(np.uint64)
    rows[::501] = np.uint64(n + 3)   # out of range: empty rows
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=6, arity=arity)
    base = gm.get_rows_csr(rows)
    small = gm.get_rows_csr(rows[:37])
    gm.close()
    monkeypatch.setenv("GGQ_EMIT_FLAT", "1")
    gf = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=6, arity=arity)
    try:
        off, cols = gf.get_rows_csr(rows)
        assert np.array_equal(off, base[0]) and np.array_equal(cols, base[1])
        off, cols = gf.get_rows_csr(rows[:37])   # ragged last tile, caller's order
        assert np.array_equal(off, small[0]) and np.array_equal(cols, small[1])
        assert np.array_equal(gf.count_labels(rows), np.bincount(base[1].astype(np.int64), minlength=m).astype(np.uint64))
    finally:
        gf.close()


def test_underestimated_stack_reruns_with_the_provable_capacity(ggq, orc, monkeypatch):
    """The shared stack capacity of the descent kernel comes from a host replay of its pop/push policy, which is an
    estimate (round-1 advisor: sparse rows of irregular trees can stack higher than the all-ones row).  The kernel
    never writes past the capacity: it raises a flag, the engine switches the matrix to the spill instantiation with
    rows-per-tile x leaf-groups capacity and reruns the batch.  GGQ_STACK_UNDERESTIMATE forces that path."""
    n, m, d = 20_000, 400, 0.35
    thr = ggq.density_threshold(d)
    rows = np.random.default_rng(4).integers(0, n, 30_000).astype(np.uint64)
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=5, arity=4)
    base = gm.get_rows_csr(rows)
    gm.close()
    monkeypatch.setenv("GGQ_STACK_UNDERESTIMATE", "12")
    for first in ("small", "large", "labels", "pipelined"):
        gs = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=5, arity=4)
        try:
            if first == "small":     # small-batch path notices the flag and falls back to the general path
                off, cols = gs.get_rows_csr(rows[:40])
                assert np.array_equal(off, base[0][:41]) and np.array_equal(cols, base[1][:int(base[0][40])])
            elif first == "labels":
                lc, lv = gs.get_labels(rows[:300], 0.0)
                exp = np.bincount(base[1][:int(base[0][300])].astype(np.int64), minlength=m)
                assert np.array_equal(lc, np.nonzero(exp)[0].astype(np.uint64)) and np.array_equal(lv, exp[exp != 0].astype(np.uint64))
            elif first == "pipelined":
                big = np.tile(rows, 80)   # 2.4e6 rows: several chunks in flight when the first one raises the flag
                off = np.zeros(big.size + 1, np.uint64)
                cols = np.zeros(int(base[0][-1]) * 80 + 16, np.uint64)
                tot = gs.get_rows_into(big, off, cols)
                assert tot == int(base[0][-1]) * 80
                assert np.array_equal(np.diff(off.astype(np.int64)), np.tile(np.diff(base[0].astype(np.int64)), 80))
                assert np.array_equal(cols[:tot], np.tile(base[1], 80))
            off, cols = gs.get_rows_csr(rows)
            assert np.array_equal(off, base[0]) and np.array_equal(cols, base[1])
            assert np.array_equal(gs.count_labels(rows), np.bincount(base[1].astype(np.int64), minlength=m).astype(np.uint64))
        finally:
            gs.close()
    for t in range(0, rows.size, 997):
        assert np.array_equal(base[1][base[0][t]:base[0][t + 1]], orc.synth_row(m, thr, 5, int(rows[t])))


@pytest.mark.parametrize("n,m,d,arity", [(50_000, 2652, 0.003, 4), (80_000, 1000, 0.004, 2), (30_000, 300, 0.01, 5),
                                         (20_000, 64, 0.3, 4)])
def test_leaf_codes_give_the_same_rows(ggq, orc, n, m, d, arity, tmp_path):
    """Leaf codes (one byte per position of a coded node's child vectors: the single label below it, or 'descend')
    are an accelerator of the descent, chosen per node at load time.  With codes on and off the engine must return
    the same rows — order included — the same histograms, and both must match direct evaluation; the image file
    carries the codes."""
    rows = np.random.default_rng(11).integers(0, n, 60_000).astype(np.uint64)
    rows[::777] = np.uint64(n + 5)
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=8, arity=arity)
    try:
        info = gm.leaf_codes
        if d < 0.05:
            assert info["nodes"] > 0 and info["in_use"] and info["bytes"] > 0, info   # sparse rows: codes must pay
        with_codes = gm.get_rows_csr(rows)
        small = gm.get_rows_csr(rows[:50])
        hist = gm.count_labels(rows)
        ws = gm.workspace()
        if info["nodes"]:
            import torch
            d_rows = torch.from_numpy(rows.view(np.int64)).cuda()
            ws.get_rows_device(d_rows.data_ptr(), rows.size, 4)
            assert ws.stat("code_probes") > 0
            assert ws.stat("code_escapes") <= ws.stat("code_probes")
        ws.close()
        gm.set_option("leaf_codes", 0)
        assert not gm.leaf_codes["in_use"]
        plain = gm.get_rows_csr(rows)
        assert np.array_equal(plain[0], with_codes[0]) and np.array_equal(plain[1], with_codes[1])
        assert np.array_equal(small[0], plain[0][:51]) and np.array_equal(small[1], plain[1][:int(plain[0][50])])
        assert np.array_equal(gm.count_labels(rows), hist)
        assert np.array_equal(hist, np.bincount(plain[1].astype(np.int64), minlength=m).astype(np.uint64))
        gm.set_option("leaf_codes", 1)
        again = gm.get_rows_csr(rows)
        assert np.array_equal(again[0], plain[0]) and np.array_equal(again[1], plain[1])
        thr = ggq.density_threshold(d)
        for t in range(0, rows.size, 499):
            exp = orc.synth_row(m, thr, 8, int(rows[t])) if rows[t] < n else np.zeros(0, np.uint64)
            assert np.array_equal(plain[1][plain[0][t]:plain[0][t + 1]], exp)
        img = str(tmp_path / "codes.ggqimg")
        gm.save_image(img)
    finally:
        gm.close()
    gi = ggq.GpuMatrix.load_image(img)
    try:
        assert gi.leaf_codes == info
        off, cols = gi.get_rows_csr(rows)
        assert np.array_equal(off, plain[0]) and np.array_equal(cols, plain[1])
    finally:
        gi.close()


@pytest.mark.parametrize("n,m,d", [(3_000_000, 2652, 0.003), (50_000, 300, 0.02), (5_000, 12_000, 0.001)])
def test_compact_host_call_matches_get_rows_into(ggq, n, m, d):
    """ggq_get_rows_compact (u32 row ids in, u8 / u16 counts and u16 / u32 ids out) returns exactly the CSR of
    ggq_get_rows_into: counts = differences of its offsets, ids bit-identical, for batches below and above the
    pipelining threshold; too-narrow count or id types are refused, never truncated."""
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=12, arity=4)
    try:
        for nq in (1, 777, 2_500_000):
            rows = np.random.default_rng(nq).integers(0, n, nq).astype(np.uint64)
            rows[::1001] = np.uint64(n + 1)
            off = np.zeros(nq + 1, np.uint64)
            cap = int(nq * (m * d * 1.5 + 4)) + 64
            cols = np.zeros(cap, np.uint64)
            tot = gm.get_rows_into(rows, off, cols)
            for rdt in (np.uint32, np.uint64):
                for cdt in (np.uint8, np.uint16, np.uint32):
                    for idt in ((np.uint16, np.uint32, np.uint64) if m <= 65536 else (np.uint32, np.uint64)):
                        if nq > 1000 and (rdt, cdt, idt) not in ((np.uint32, np.uint8, np.uint16 if m <= 65536 else np.uint32),
                                                                 (np.uint64, np.uint32, np.uint64)):
                            continue   # every combination on the small batches, two on the large one
                        cnt = np.zeros(nq, cdt)
                        ids = np.zeros(cap, idt)
                        r = rows.astype(rdt) if rdt == np.uint64 or n + 1 < 2**32 else rows
                        got = gm.get_rows_compact(r, cnt, ids)
                        assert got == tot
                        assert np.array_equal(cnt.astype(np.int64), np.diff(off.astype(np.int64)))
                        assert np.array_equal(ids[:tot].astype(np.uint64), cols[:tot])
        if m > 65536:
            with pytest.raises(ggq.GgqError):
                gm.get_rows_compact(rows[:10].astype(np.uint32), np.zeros(10, np.uint8), np.zeros(100, np.uint16))
    finally:
        gm.close()
    # a row with more ids than a u8 count can hold: GGQ_ERR_LIMIT, and it works with u16 counts
    gd = ggq.GpuMatrix.synthesize_brwt(2_000, 600, 0.9, seed=1, arity=4)
    try:
        rows = np.arange(100, dtype=np.uint32)
        with pytest.raises(ggq.GgqError) as e:
            gd.get_rows_compact(rows, np.zeros(100, np.uint8), np.zeros(100 * 600, np.uint16))
        assert e.value.code == -5   # GGQ_ERR_LIMIT
        cnt = np.zeros(100, np.uint16)
        ids = np.zeros(100 * 600, np.uint16)
        tot = gd.get_rows_compact(rows, cnt, ids)
        o, c = gd.get_rows_csr(rows.astype(np.uint64))
        assert tot == c.size and np.array_equal(cnt, np.diff(o.astype(np.int64)).astype(np.uint16)) and np.array_equal(ids[:tot], c.astype(np.uint16))
    finally:
        gd.close()


def test_pipelined_calls_on_the_c4_column_geometry_against_direct_evaluation(ggq):
    """A batch above 2^20 rows (the chunk pipeline of ggq_get_rows_into / ggq_get_rows_compact / ggq_count_labels) on
    the benchmark's tile geometry (2652 columns: 83 bitmap words, 3 summary words per row, leaf codes in use), checked
    against an expectation that owes nothing to the engine: direct evaluation of the implicit matrix
    M[i][j] = mix64(seed*K + i*m + j) < T on a 1-in-23 sample of the batch, and the histogram of that evaluation."""
    import bench
    n, m, d = 4_000_000, 2652, 0.003
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=bench.SEED, arity=4)
    try:
        assert gm.leaf_codes["in_use"]
        rows = np.random.default_rng(8).integers(0, n, (1 << 21) + 54_321).astype(np.uint64)
        off = np.empty(rows.size + 1, np.uint64)
        cols = np.empty(int(rows.size * m * d * 1.2) + 4096, np.uint32)
        total = gm.get_rows_into(rows, off, cols)
        pick = np.arange(0, rows.size, 23)
        counts, ids = bench.direct_rows(rows[pick], m, d, threads=8)
        assert np.array_equal(np.diff(off.astype(np.int64))[pick], counts.astype(np.int64))
        starts = off[pick].astype(np.int64)
        got = np.concatenate([cols[s:s + c] for s, c in zip(starts, counts.astype(np.int64))])
        assert np.array_equal(got.astype(np.uint64), ids)
        # the compact call on the same batch
        cnt8 = np.empty(rows.size, np.uint8)
        ids16 = np.empty(cols.size, np.uint16)
        assert gm.get_rows_compact(rows.astype(np.uint32), cnt8, ids16) == total
        assert np.array_equal(cnt8.astype(np.int64), np.diff(off.astype(np.int64)))
        assert np.array_equal(ids16[:total], cols[:total].astype(np.uint16))
        # count_labels over a batch above 2^22 rows (its own chunk pipeline): histogram of the direct evaluation,
        # the sample taken four times
        big = np.tile(rows[pick], 52)[: (1 << 22) + 1000]
        reps = np.bincount(np.arange(big.size) % pick.size, minlength=pick.size)
        hist = np.bincount(ids.astype(np.int64), weights=np.repeat(reps, counts.astype(np.int64)), minlength=m)
        assert np.array_equal(gm.count_labels(big), hist.astype(np.uint64))
    finally:
        gm.close()


def test_label_batch_forms_alternate_on_one_matrix(ggq):
    """The direct (<= 2^14 indices) and the general form of the label batch share a workspace's scratch: alternate
    them on one matrix (a direct call first sized the shared input buffer and starved the general form's buffers
    once) and compare every answer with the single-query call."""
    n, m, d = 200_000, 300, 0.02
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=3, arity=4)
    try:
        rng = np.random.default_rng(2)
        for sizes in ([300] * 5, [300] * 80, [7, 300, 1], [250] * 100, [300] * 3):
            qoff = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
            rows = rng.integers(0, n, int(qoff[-1])).astype(np.uint64)
            ratios = rng.choice([0.0, 0.05, 0.5], size=len(sizes))
            loff, cols, cnt = gm.get_labels_batch(rows, qoff, ratios)
            for q in range(0, len(sizes), max(1, len(sizes) // 7)):
                ec, ev = gm.get_labels(rows[int(qoff[q]):int(qoff[q + 1])], float(ratios[q]))
                a, b = int(loff[q]), int(loff[q + 1])
                assert np.array_equal(cols[a:b], ec) and np.array_equal(cnt[a:b], ev), (len(sizes), q)
    finally:
        gm.close()
