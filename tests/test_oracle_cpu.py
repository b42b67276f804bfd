"""CPU suite (no GPU): pins the oracle.

* the plain-C restatement (oracle/brwt_oracle.c) against the committed golden fixtures — the reference's
  own serialized streams and the reference's own answers (tests/golden/make_golden.py);
* the restatement against the live reference (oracle/_ref/libggref.so) on seeded random inputs, when the
  reference build is present;
* the reference's own gtest files for the path (tests/test_{BRWT,BRWT_optimizer,bit_vector,column_major}.cpp)
  run against the oracle build, when /root/reference is present.
"""
import glob
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, random_dense

GOLD = os.path.join(ROOT, "tests", "golden")


def load_gold(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


MATS = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLD, "brwt_*.npz")))
BVS = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLD, "bv_*.npz")))


def test_fixtures_exist():
    assert len(MATS) >= 6 and len(BVS) >= 5


@pytest.mark.parametrize("name", MATS)
def test_c_oracle_brwt_vs_golden(orc, name):
    g = load_gold(name)
    data = g["stream"].tobytes()
    m = orc.OrcBRWT(data)
    assert m.consumed == len(data)
    n, ncol, rel, nodes, tcs = [int(x) for x in g["stats"]]
    assert (m.num_rows, m.num_columns, m.num_relations, m.num_nodes, m.total_column_size) == (n, ncol, rel, nodes, tcs)
    off, cols = g["row_offsets"], g["row_cols"]
    shape = tuple(int(x) for x in g["shape"])
    dense = np.unpackbits(g["dense"])[: shape[0] * shape[1]].reshape(shape) if shape[0] * shape[1] else np.zeros(shape, np.uint8)
    touches = 0
    for r in range(n):
        got = m.get_row(r)
        assert np.array_equal(got, cols[off[r]:off[r + 1]]), r          # order included
        assert set(got.tolist()) == set(np.nonzero(dense[r])[0].tolist())
        touches += m.row_touches(r)
    assert touches == tcs   # SURVEY §6 identity: every bit of every node vector is visited exactly once
    coff, crows = g["col_offsets"], g["col_rows"]
    for c in range(ncol):
        assert np.array_equal(m.get_column(c), crows[coff[c]:coff[c + 1]]), c
        for r in range(0, n, 7):
            assert m.get(r, c) == bool(dense[r, c])


def test_c_oracle_colmajor_vs_golden(orc):
    g = load_gold("colmajor_200x30")
    data = g["stream"].tobytes()
    m = orc.OrcColMajor(data)
    n, ncol, rel = [int(x) for x in g["stats"]]
    assert (m.consumed, m.num_rows, m.num_columns, m.num_relations) == (len(data), n, ncol, rel)
    off, cols = g["row_offsets"], g["row_cols"]
    for r in range(n):
        assert np.array_equal(m.get_row(r), cols[off[r]:off[r + 1]])
    coff, crows = g["col_offsets"], g["col_rows"]
    for c in range(ncol):
        assert np.array_equal(m.get_column(c), crows[coff[c]:coff[c + 1]])


@pytest.mark.parametrize("name", BVS)
@pytest.mark.parametrize("kind", ["stat", "sd", "rrr"])
def test_c_oracle_bitvector_vs_golden(orc, name, kind):
    g = load_gold(name)
    n = int(g["n"][0])
    data = g["stream_" + kind].tobytes()
    v = orc.OrcBits.load(kind, data)
    assert v.consumed == len(data) and v.size == n
    pos = np.arange(n + 3, dtype=np.uint64)
    assert np.array_equal(v.rank1(pos), g["rank1_" + kind])
    if n:
        assert np.array_equal(v.get(pos[:n]), g["get_" + kind])
    sel = g["select1_" + kind]
    if sel.size:
        assert np.array_equal(v.select1(np.arange(1, sel.size + 1)), sel)
    assert v.num_set_bits == sel.size
    if kind == "sd":
        assert int(v.is_inverted) == int(g["inverted_sd"][0])


def test_golden16_known_answers(orc):
    # tests/test_bit_vector.cpp:88-94: the 16-bit pattern and its hand-checkable answers
    bits = [0, 1, 0, 1, 1, 1, 1, 0, 0, 1, 0, 0, 0, 0, 1, 1]
    g = load_gold("bv_golden16")
    for kind in ("stat", "sd", "rrr"):
        assert g["rank1_" + kind][:16].tolist() == np.cumsum(bits).tolist()
        assert g["rank1_" + kind][16:].tolist() == [8, 8, 8]          # clamped beyond the end
        assert g["select1_" + kind].tolist() == [i for i, b in enumerate(bits) if b]


@pytest.mark.parametrize("n,m,d,kw", [
    (500, 17, 0.3, dict(arity=2)), (2000, 64, 0.05, dict(arity=4, greedy=True, relax=5)),
    (300, 1, 0.5, dict(arity=2)), (100, 7, 1.0, dict(arity=3)), (100, 7, 0.0, dict(arity=3)),
    (1500, 130, 0.02, dict(arity=2, greedy=True, relax=9)),
])
def test_c_oracle_vs_live_reference(orc, ref, n, m, d, kw):
    dense = random_dense(n, m, d, seed=n + m)
    rm = ref.RefMatrix.build_brwt(ref.pack_columns(dense), n, **kw)
    data = rm.serialize()
    om = orc.OrcBRWT(data)
    assert (om.num_nodes, om.total_column_size, om.num_relations) == (rm.num_nodes, rm.total_column_size, rm.num_relations)
    for r in range(0, n, max(1, n // 300)):
        assert np.array_equal(om.get_row(r), rm.get_row(r))
    for c in range(m):
        assert np.array_equal(om.get_column(c), rm.get_column(c))


def test_synth_generator_agrees(orc, ref, ggq):
    # the portable implicit matrix: C oracle row == reference-side generator == numpy mix64
    n, m, d = 300, 50, 0.1
    thr = ggq.density_threshold(d)
    cw = ref.gen_implicit(n, m, thr, 42)
    dense = ref.unpack_columns(cw, n)
    for r in (0, 1, 150, 299):
        assert np.array_equal(orc.synth_row(m, thr, 42, r), np.nonzero(dense[r])[0].astype(np.uint64))
    with np.errstate(over="ignore"):
        x = np.uint64((42 * ggq.SEED_MUL) % 2**64) + np.uint64(5) * np.uint64(m) + np.arange(m, dtype=np.uint64)
    assert np.array_equal((ggq.mix64(x) < np.uint64(thr)).astype(np.uint8), dense[5])
    assert orc.mix64(12345) == int(ggq.mix64(np.array([12345], dtype=np.uint64))[0])


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="reference sources absent")
def test_reference_unit_tests_pass_on_oracle_build():
    """The reference's own 94 gtest cases for the path, compiled against the oracle build."""
    odir = os.path.join(ROOT, "oracle")
    p = subprocess.run(["make", "-C", odir, "-j8", "ref_tests"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert p.returncode == 0, p.stdout[-3000:]
    run = os.path.join(odir, "_ref", "run")
    os.makedirs(run, exist_ok=True)
    p = subprocess.run(["../ref_unit_tests"], cwd=run, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                       text=True, timeout=900)
    tail = p.stdout[-2000:]
    assert p.returncode == 0, tail
    assert "PASSED" in tail and "FAILED" not in tail, tail


def test_space_model_constant_of_the_device_builder(ref):
    """ggq_build_greedy.cuh restates BRWTOptimizer's RRR space model with sizeof(bit_vector_rrr<>) as a constant."""
    import ctypes
    import re
    f = ref.lib().ggref_sizeof_bit_vector_rrr
    f.restype = ctypes.c_uint64
    src = open(os.path.join(ROOT, "genome_graph_annotation_b200", "csrc", "ggq_build_greedy.cuh")).read()
    k = re.search(r"kSizeofBitVectorRrrBits = (\d+) \* 8", src)
    assert k and int(k.group(1)) == f()


def test_threaded_rrr_decode_of_large_vectors(ref, monkeypatch):
    """Vectors of >= 2^12 superblocks are decoded by several host threads (ggq_format.hpp: load_rrr63); the parse of
    a 9e6-row reference stream must give the reference's statistics with 1, 3 and many threads, and a stream whose
    payload is damaged must still be rejected, not crash."""
    import genome_graph_annotation_b200 as g
    n = 9_000_000
    cw = ref.gen_norepl(n, 3, 0.2, seed=3)
    rm = ref.RefMatrix.build_brwt(cw, n, arity=2, threads=8)
    data = rm.serialize()
    for thr in ("1", "3", "16"):
        monkeypatch.setenv("GGQ_LOAD_THREADS", thr)
        info = g.inspect_brwt(data)
        assert (info["num_rows"], info["num_columns"], info["num_relations"], info["num_nodes"], info["total_column_size"]) == \
            (n, 3, rm.num_relations, rm.num_nodes, rm.total_column_size), thr
        assert info["consumed"] == len(data)
    bad = bytearray(data)
    for i in range(len(bad) // 2, len(bad) // 2 + 4096):
        bad[i] = 0xFF
    try:
        info = g.inspect_brwt(bytes(bad))   # flipped offsets may still decode to SOME vector; it must not crash
        assert info["num_rows"] == n
    except g.GgqError as e:
        assert e.code == -1


@pytest.mark.parametrize("n", [63, 126, 2015, 2016, 2017, 4032, 2016 * 5, 2016 * 100, 2016 * 4096, 2016 * 4096 + 63])
@pytest.mark.parametrize("lazy", [True, False])
def test_vector_lengths_at_the_rrr_block_boundaries(ref, monkeypatch, n, lazy):
    """sdsl appends an empty dummy block when the length is a multiple of 63; when that block opens a superblock (length
    a multiple of 2016 = 32 x 63) the superblock's offset pointer and rank sample are 0.  Such a stream is valid and must
    parse to the reference's statistics (found by the 7e8-row load probe: its root vector is 2016 x 347222 long)."""
    import genome_graph_annotation_b200 as g
    if not lazy:
        monkeypatch.setenv("GGQ_LOAD_EAGER", "1")
    dense = random_dense(n, 3, 0.3, n % 1000) if n < 10 ** 6 else None
    cw = ref.pack_columns(dense) if dense is not None else ref.gen_norepl(n, 3, 0.3, seed=5)
    rm = ref.RefMatrix.build_brwt(cw, n, arity=2, threads=4)
    data = rm.serialize()
    info = g.inspect_brwt(data)
    assert (info["num_rows"], info["num_relations"], info["num_nodes"], info["total_column_size"]) == \
        (n, rm.num_relations, rm.num_nodes, rm.total_column_size)
    assert info["consumed"] == len(data)


def _pack_bits(bits):
    n = bits.size
    pad = np.zeros(((n + 63) // 64) * 64, np.uint8)
    pad[:n] = bits
    return np.packbits(pad, bitorder="little").view(np.uint64).copy() if n else np.zeros(0, np.uint64)


def test_rrr_decoder_reproduces_the_bits_of_reference_streams(ref, monkeypatch):
    """The loader's rrr_vector<63> decoder (ggq_rrr63_decode_buffer = read_rrr63_raw + decode_rrr63: eight blocks walked
    together, branch-free) against streams serialized by the REFERENCE's bit_vector_rrr on known bits: every bit, at
    every density (all classes 0..63 occur, both sides of the inversion rule), at the lengths where sdsl's layout has
    corners, single- and multi-threaded; damaged offsets are an error or some other vector, never a crash."""
    import genome_graph_annotation_b200 as g
    rng = np.random.default_rng(7)
    for n in [0, 1, 5, 62, 63, 64, 125, 126, 127, 2015, 2016, 2017, 4032, 63 * 32 * 3 + 1, 100_003, 1 << 19]:
        for d in [0.0, 0.001, 0.02, 0.1, 0.3, 0.5, 0.7, 0.9, 0.99, 1.0]:
            bits = (rng.random(n) < d).astype(np.uint8)
            w = _pack_bits(bits)
            data = ref.RefBitVector.build("rrr", w, n).serialize()
            got, n_bits, used = g.rrr63_decode(data)
            assert n_bits == n and used == len(data), (n, d)
            assert np.array_equal(got, w), (n, d)
    # every class in one vector: block b has (b mod 64) ones at random places
    blocks = []
    for b in range(64 * 40):
        x = np.zeros(63, np.uint8)
        x[rng.permutation(63)[: b % 64]] = 1
        blocks.append(x)
    bits = np.concatenate(blocks)
    w = _pack_bits(bits)
    got, n_bits, _ = g.rrr63_decode(ref.RefBitVector.build("rrr", w, bits.size).serialize())
    assert n_bits == bits.size and np.array_equal(got, w)
    # long vectors: several host threads, dense and sparse stretches
    for thr in ("1", "5"):
        monkeypatch.setenv("GGQ_LOAD_THREADS", thr)
        n = 2016 * 5000 + 17
        bits = (rng.random(n) < 0.25).astype(np.uint8)
        bits[n // 3:n // 2] = (rng.random(n // 2 - n // 3) < 0.95).astype(np.uint8)
        w = _pack_bits(bits)
        data = ref.RefBitVector.build("rrr", w, n).serialize()
        got, n_bits, _ = g.rrr63_decode(data)
        assert n_bits == n and np.array_equal(got, w), thr
    bad = bytearray(data)
    for i in range(len(bad) // 2, len(bad) // 2 + 512):
        bad[i] ^= 0xFF
    try:
        got, n_bits, _ = g.rrr63_decode(bytes(bad))
        assert n_bits == n
    except g.GgqError as e:
        assert e.code == -1


def test_inspect_sizes_the_image_in_the_selected_slot_layout(ref):
    """Host-only deployment sizing (ggq_brwt_inspect_buffer, no GPU): the capacity layout (32-byte slots of 224
    positions) stores 8/7 of the plain bits instead of 4/3 — same tree, same statistics, smaller image;
    both sizes bracketed by the plain bits of the reference's own total_column_size."""
    import genome_graph_annotation_b200 as g
    n = 2_000_000
    cw = ref.gen_norepl(n, 16, 0.2, seed=9)
    rm = ref.RefMatrix.build_brwt(cw, n, arity=4, threads=8)   # (arity 4: every sibling group fills its four lanes)
    data = rm.serialize()
    try:
        g.set_default_option("wide_slots", 0)
        a = g.inspect_brwt(data)
        g.set_default_option("wide_slots", 1)
        b = g.inspect_brwt(data)
    finally:
        g.set_default_option("wide_slots", 0)
    for k in ("num_rows", "num_columns", "num_nodes", "num_relations", "total_column_size", "groups", "max_depth"):
        assert a[k] == b[k], k
    assert np.array_equal(a["dfs_columns"], b["dfs_columns"])
    plain = rm.total_column_size / 8
    assert a["total_column_size"] == rm.total_column_size
    assert 4 / 3 * plain <= a["device_bytes"] < 1.40 * plain      # rank header in every 16-byte slot (+ samples, padding)
    assert 8 / 7 * plain <= b["device_bytes"] < 1.20 * plain      # one header per 32-byte slot
    with pytest.raises(g.GgqError):
        g.set_default_option("no_such_option", 1)


def test_rrr_writer_is_byte_identical_to_sdsl(ref, monkeypatch):
    """ggq_rrr63_encode_buffer (host only; the building block of ggq_brwt_serialize) against the reference's
    bit_vector_rrr<63>::serialize on the same bits: every array of sdsl's rrr_vector — block classes, offsets in
    sdsl's numbering, superblock pointers, rank samples, inversion flags, widths and padding — byte for byte, over
    the lengths where the construction rule has corners (multiples of 63 and of 2016, partial last blocks) and
    densities on both sides of the inversion rule."""
    import genome_graph_annotation_b200 as g
    rng = np.random.default_rng(1)
    for n in [0, 1, 5, 62, 63, 64, 125, 126, 127, 2015, 2016, 2017, 4031, 4032, 4033, 63 * 32 * 3, 100_003, 1 << 19]:
        for d in [0.0, 0.001, 0.1, 0.5, 0.6, 0.9, 1.0]:
            bits = (rng.random(n) < d).astype(np.uint8)
            pad = np.zeros(((n + 63) // 64) * 64, np.uint8)
            pad[:n] = bits
            w = np.packbits(pad, bitorder="little").view(np.uint64).copy() if n else np.zeros(0, np.uint64)
            assert g.rrr63_encode(w, n) == ref.RefBitVector.build("rrr", w, n).serialize(), (n, d)
    # a run of dense superblocks followed by sparse ones: inversion switches inside the vector
    bits = np.concatenate([np.ones(2016 * 3, np.uint8), (rng.random(2016 * 2 + 100) < 0.02).astype(np.uint8),
                           (rng.random(5000) < 0.8).astype(np.uint8)])
    n = bits.size
    pad = np.zeros(((n + 63) // 64) * 64, np.uint8)
    pad[:n] = bits
    w = np.packbits(pad, bitorder="little").view(np.uint64).copy()
    assert g.rrr63_encode(w, n) == ref.RefBitVector.build("rrr", w, n).serialize()
    # long vectors are encoded by several host threads (>= 2^16 blocks): dummy-block and partial-block endings, with
    # dense and sparse stretches so that the inversion flag flips along the vector
    monkeypatch.setenv("GGQ_LOAD_THREADS", "5")
    for n in (63 * 65536, 63 * 65536 + 1, 2016 * 2100 + 62):
        bits = (rng.random(n) < 0.02).astype(np.uint8)
        bits[n // 3:n // 2] = (rng.random(n // 2 - n // 3) < 0.9).astype(np.uint8)
        pad = np.zeros(((n + 63) // 64) * 64, np.uint8)
        pad[:n] = bits
        w = np.packbits(pad, bitorder="little").view(np.uint64).copy()
        assert g.rrr63_encode(w, n) == ref.RefBitVector.build("rrr", w, n).serialize(), n
