"""GPU parity of bit_vector_{stat,sd,rrr} rank1 / select1 / get (K5) and rank0 / select0 / get_int / to_vector /
call_ones through the C ABI vs the unmodified
reference classes (common/bit_vector.cpp:281-304, 690-718, 857-888 behind oracle/_ref/libggref.so).
Cases follow tests/test_bit_vector.cpp: the 16-bit golden pattern (:88-89), all-0 / all-1 vectors incl.
rank1 far beyond the end (:58-86), rank clamping (:25-29), the sd inversion rule over all 10-bit words
(:642-658), serialized round trips (:307-461), and the SURVEY §4 length/density grid."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

KINDS = ["stat", "sd", "rrr"]
GOLDEN16 = [0, 1, 0, 1, 1, 1, 1, 0, 0, 1, 0, 0, 0, 0, 1, 1]


def pack(bits):
    bits = np.asarray(bits, dtype=np.uint8)
    n = bits.size
    pad = np.zeros(((n + 63) // 64) * 64, dtype=np.uint8)
    pad[:n] = bits
    return np.packbits(pad, bitorder="little").view(np.uint64).copy() if n else np.zeros(0, np.uint64)


def check(ggq, ref, kind, bits, via_stream=False):
    bits = np.asarray(bits, dtype=np.uint8)
    n = bits.size
    words = pack(bits)
    rv = ref.RefBitVector.build(kind, words, n)
    if via_stream:
        data = rv.serialize()
        gv = ggq.GpuBitVector.load(kind, data)
        assert gv.consumed == len(data)
    else:
        gv = ggq.GpuBitVector.build(kind, words, n)
    try:
        ones = int(bits.sum())
        assert gv.size == n == rv.size
        assert gv.num_set_bits == ones == rv.num_set_bits
        if kind == "sd":
            assert gv.is_inverted == rv.is_inverted == (ones > n // 2)
        if n <= 5000:
            pos = np.arange(n, dtype=np.uint64)
        else:
            rng = np.random.default_rng(n)
            pos = np.unique(np.concatenate([rng.integers(0, n, 4000), [0, 1, 62, 63, 64, n - 1, n - 2, n // 2]])).astype(np.uint64)
        # rank: inside, at and beyond the end (clamped)
        rq = np.concatenate([pos, np.array([n, n + 1, n + 990, 2**40, 2**64 - 1], dtype=np.uint64)])
        assert np.array_equal(gv.rank1(rq), rv.rank1(rq))
        if n:
            assert np.array_equal(gv.get(pos), rv.get(pos).astype(np.uint8))
            assert np.array_equal(gv.get(pos), bits[pos.astype(np.int64)])
        if ones:
            ks = np.arange(1, ones + 1, dtype=np.uint64) if ones <= 5000 else \
                np.unique(np.concatenate([np.random.default_rng(1).integers(1, ones + 1, 4000), [1, 2, ones]])).astype(np.uint64)
            sel = gv.select1(ks)
            assert np.array_equal(sel, rv.select1(ks))
            assert np.array_equal(sel, np.nonzero(bits)[0][ks.astype(np.int64) - 1].astype(np.uint64))
        # the rest of the bit_vector interface: rank0, select0, get_int, to_vector, call_ones
        rq0 = np.concatenate([pos, np.array([n, n + 1, n + 990, 2**40], dtype=np.uint64)])
        assert np.array_equal(gv.rank0(rq0), rv.rank0(rq0))
        zeros = n - ones
        if zeros:
            kz = np.arange(1, zeros + 1, dtype=np.uint64) if zeros <= 5000 else \
                np.unique(np.concatenate([np.random.default_rng(2).integers(1, zeros + 1, 4000), [1, 2, zeros]])).astype(np.uint64)
            sel0 = gv.select0(kz)
            assert np.array_equal(sel0, np.nonzero(bits == 0)[0][kz.astype(np.int64) - 1].astype(np.uint64))
            if kind == "rrr":   # the only kind the reference offers select0 on (bit_vector.hpp:199)
                assert np.array_equal(sel0, rv.select0(kz))
        for width in (1, 7, 32, 64):
            if n >= width:
                pq = pos[pos + np.uint64(width) <= np.uint64(n)][:2000]
                assert np.array_equal(gv.get_int(pq, width), rv.get_int(pq, width)), (kind, width)
        assert np.array_equal(gv.to_words(), words[:(n + 63) // 64])
        mine = gv.call_ones()
        assert np.array_equal(mine, np.nonzero(bits)[0].astype(np.uint64))
        assert np.array_equal(np.sort(rv.call_ones()), mine)
    finally:
        gv.close()


@pytest.mark.parametrize("kind", KINDS)
def test_golden_and_degenerate(ggq, ref, kind):
    check(ggq, ref, kind, GOLDEN16)
    check(ggq, ref, kind, np.zeros(10, np.uint8))
    check(ggq, ref, kind, np.ones(10, np.uint8))
    check(ggq, ref, kind, np.zeros(0, np.uint8))
    check(ggq, ref, kind, [1])
    check(ggq, ref, kind, [0])


def test_sd_inversion_all_10bit_words(ggq, ref):
    for w in range(1024):   # every word the reference pins (tests/test_bit_vector.cpp:642-658)
        bits = [(w >> i) & 1 for i in range(10)]
        check(ggq, ref, "sd", bits)


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("n", [1, 62, 63, 64, 447, 448, 449, 1007, 1008, 1009, 2015, 2016, 2017, 2**20 + 1])
def test_length_density_grid(ggq, ref, kind, n):
    rng = np.random.default_rng(n * 3 + len(kind))
    for d in (0.0, 1e-3, 0.3, 0.5, 0.7, 1.0):
        bits = (rng.random(n) < d).astype(np.uint8)
        check(ggq, ref, kind, bits)


@pytest.mark.parametrize("kind", KINDS)
def test_serialized_stream_roundtrip(ggq, ref, kind):
    rng = np.random.default_rng(11)
    for n, d in ((1000, 0.1), (5000, 0.9), (70000, 0.01), (0, 0.0)):
        check(ggq, ref, kind, (rng.random(n) < d).astype(np.uint8), via_stream=True)


def test_bare_sd_vector_stream_is_rejected(ggq, ref):
    # tests/test_bit_vector.cpp:432-461: an sd_vector without the trailing `inverted` byte must not load
    bits = (np.random.default_rng(2).random(1000) < 0.1).astype(np.uint8)
    data = ref.RefBitVector.build("sd", pack(bits), 1000).serialize()
    with pytest.raises(ggq.GgqError):
        ggq.GpuBitVector.load("sd", data[:-1])


@pytest.mark.parametrize("kind", KINDS)
def test_clustered_and_skewed(ggq, ref, kind):
    n = 300_000
    bits = np.zeros(n, np.uint8)
    bits[1000:1500] = 1
    bits[250_000:250_003] = 1
    bits[n - 1] = 1
    check(ggq, ref, kind, bits)
    bits = np.ones(n, np.uint8)
    bits[5000:90000] = 0
    check(ggq, ref, kind, bits)


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("d", [0.001, 0.1, 0.5])
def test_synthesized_matches_oracle_generator(ggq, orc, ref, kind, d):
    # config C3 generator at 2^22 bits: bit i = mix64(42*K + i) < T, same bits from the C oracle
    n = 1 << 22
    thr = ggq.density_threshold(d)
    words = orc.synth_bits(n, thr, 42)
    gv = ggq.GpuBitVector.synthesize(kind, n, d, seed=42)
    rv = ref.RefBitVector.build(kind, words, n)
    assert gv.num_set_bits == rv.num_set_bits
    rng = np.random.default_rng(3)
    pos = rng.integers(0, n, 20000).astype(np.uint64)
    assert np.array_equal(gv.rank1(pos), rv.rank1(pos))
    assert np.array_equal(gv.get(pos), rv.get(pos).astype(np.uint8))
    ks = rng.integers(1, rv.num_set_bits + 1, 20000).astype(np.uint64)
    assert np.array_equal(gv.select1(ks), rv.select1(ks))
    gv.close()


@pytest.mark.parametrize("kind", KINDS)
def test_map_rank1_is_graph_to_anno_index(ggq, ref, kind):
    """ggq_map_rank1 = AnnotatedDBG::graph_to_anno_index over a batch (common/annotated_dbg.cpp:142-156):
    annotation_mask.rank1(i) - 1 where the mask is set; unset / out-of-range indices come back as UINT64_MAX."""
    n = 100_000
    bits = (np.random.default_rng(3).random(n) < 0.4).astype(np.uint8)
    words = pack(bits)
    gv = ggq.GpuBitVector.build(kind, words, n)
    rv = ref.RefBitVector.build(kind, words, n)
    try:
        q = np.concatenate([np.random.default_rng(4).integers(0, n, 20_000), [0, n - 1, n, n + 5, 2**50]]).astype(np.uint64)
        got, bad = gv.map_rank1(q)
        inside = q < n
        set_ = np.zeros(q.size, bool)
        set_[inside] = bits[q[inside].astype(np.int64)] == 1
        exp = np.full(q.size, 2**64 - 1, dtype=np.uint64)
        exp[set_] = rv.rank1(q[set_]) - np.uint64(1)
        assert np.array_equal(got, exp) and bad == int((~set_).sum())
    finally:
        gv.close()
