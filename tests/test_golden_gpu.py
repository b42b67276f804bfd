"""GPU parity against the committed golden fixtures (tests/golden/*.npz: the reference's serialized streams
and the reference's own answers) — needs neither /root/reference nor oracle/_ref at run time."""
import glob
import os

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "brwt_*.npz")) + glob.glob(os.path.join(GOLD, "colmajor_*.npz"))))
def test_matrix_fixture(ggq, path):
    g = np.load(path)
    data = g["stream"].tobytes()
    kind = str(g["kind"])
    gm = ggq.GpuMatrix.load_brwt(data) if kind == "brwt" else ggq.GpuMatrix.load_colmajor(data)
    st = [int(x) for x in g["stats"]]
    assert (gm.num_rows, gm.num_columns, gm.num_relations) == tuple(st[:3])
    if kind == "brwt":
        assert (gm.num_nodes, gm.total_column_size) == tuple(st[3:5])
    n, m = st[0], st[1]
    off, cols = gm.get_rows_csr(np.arange(n, dtype=np.uint64))
    assert np.array_equal(off, g["row_offsets"]) and np.array_equal(cols, g["row_cols"])
    coff, crows = g["col_offsets"], g["col_rows"]
    buf = np.empty(max(n, 1), np.uint64)
    for c in range(m):
        assert np.array_equal(gm.get_column(c), crows[coff[c]:coff[c + 1]])
        k = gm.get_column_into(c, buf)   # caller-provided buffer; the count is a host-side lookup
        assert k == gm.get_column_count(c) == coff[c + 1] - coff[c]
        assert np.array_equal(buf[:k], crows[coff[c]:coff[c + 1]])
    if m and coff[1] - coff[0] > 0:
        with pytest.raises(ggq.GgqError):
            gm.get_column_into(0, np.empty(0, np.uint64))   # capacity too small
    shape = tuple(int(x) for x in g["shape"])
    if n * m:
        dense = np.unpackbits(g["dense"])[: n * m].reshape(shape)
        rr, cc = np.meshgrid(np.arange(n, dtype=np.uint64), np.arange(m, dtype=np.uint64), indexing="ij")
        assert np.array_equal(gm.get_cells(rr.ravel(), cc.ravel()).reshape(shape), dense)
    gm.close()


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "bv_*.npz"))))
@pytest.mark.parametrize("kind", ["stat", "sd", "rrr"])
def test_bitvector_fixture(ggq, path, kind):
    g = np.load(path)
    n = int(g["n"][0])
    gv = ggq.GpuBitVector.load(kind, g["stream_" + kind].tobytes())
    pos = np.arange(n + 3, dtype=np.uint64)
    assert gv.size == n
    assert np.array_equal(gv.rank1(pos), g["rank1_" + kind])
    if n:
        assert np.array_equal(gv.get(pos[:n]), g["get_" + kind].astype(np.uint8))
    sel = g["select1_" + kind]
    assert gv.num_set_bits == sel.size
    if sel.size:
        assert np.array_equal(gv.select1(np.arange(1, sel.size + 1, dtype=np.uint64)), sel)
    if kind == "sd":
        assert int(gv.is_inverted) == int(g["inverted_sd"][0])
    gv.close()


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "brwt_*.npz"))))
def test_image_file_round_trip(ggq, path, tmp_path):
    """Flattened device image (ggq_image_save / ggq_image_load): a matrix restored from the image
    answers exactly like the reference — rows (order included), columns, cells, counters, stats."""
    g = np.load(path)
    gm = ggq.GpuMatrix.load_brwt(g["stream"].tobytes())
    img = str(tmp_path / "m.ggqimg")
    gm.save_image(img)
    gm.close()
    gi = ggq.GpuMatrix.load_image(img)
    try:
        # the image carries the host tree: a restored matrix serializes back to the reference's stream byte for byte
        assert gi.serialize() == g["stream"].tobytes()
        st = [int(x) for x in g["stats"]]
        assert (gi.num_rows, gi.num_columns, gi.num_relations, gi.num_nodes, gi.total_column_size) == tuple(st[:5])
        n, m = st[0], st[1]
        off, cols = gi.get_rows_csr(np.arange(n, dtype=np.uint64))
        assert np.array_equal(off, g["row_offsets"]) and np.array_equal(cols, g["row_cols"])
        coff, crows = g["col_offsets"], g["col_rows"]
        for c in range(m):
            assert np.array_equal(gi.get_column(c), crows[coff[c]:coff[c + 1]])
        if n * m:
            shape = tuple(int(x) for x in g["shape"])
            dense = np.unpackbits(g["dense"])[: n * m].reshape(shape)
            rr, cc = np.meshgrid(np.arange(n, dtype=np.uint64), np.arange(m, dtype=np.uint64), indexing="ij")
            assert np.array_equal(gi.get_cells(rr.ravel(), cc.ravel()).reshape(shape), dense)
            assert np.array_equal(gi.count_labels(np.arange(n, dtype=np.uint64)), dense.sum(0).astype(np.uint64))
    finally:
        gi.close()
    # damaged files are rejected like a bad stream (GGQ_ERR_FORMAT), never half-loaded
    raw = open(img, "rb").read()
    for bad in (raw[: len(raw) // 2], b"GGQIMG99" + raw[8:], raw[:200] + bytes([raw[200] ^ 0xFF]) + raw[201:]):
        open(img, "wb").write(bad)
        with pytest.raises(ggq.GgqError) as e:
            ggq.GpuMatrix.load_image(img)
        assert e.value.code == -1


def test_image_file_of_a_device_built_shard(ggq, tmp_path):
    """Images also carry device-built trees and the shard plan (owned columns, remote lanes)."""
    n, m, d = 30_000, 90, 0.05
    rows = np.random.default_rng(2).integers(0, n, 5000).astype(np.uint64)
    for shard, n_shards in ((0, 1), (1, 3)):
        if n_shards == 1:
            gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=3, arity=4)
        else:
            gm = ggq.GpuMatrix.synthesize_brwt_shard(n, m, d, shard, n_shards, seed=3, arity=4)
        exp = gm.get_rows_csr(rows)
        stats = (gm.num_rows, gm.num_columns, gm.num_relations, gm.num_nodes, gm.total_column_size)
        img = str(tmp_path / ("s%d.ggqimg" % n_shards))
        gm.save_image(img)
        gm.close()
        gi = ggq.GpuMatrix.load_image(img)
        try:
            assert (gi.num_rows, gi.num_columns, gi.num_relations, gi.num_nodes, gi.total_column_size) == stats
            off, cols = gi.get_rows_csr(rows)
            assert np.array_equal(off, exp[0]) and np.array_equal(cols, exp[1])
            owned = [c for c in range(m) if (cols == c).any()][:3]
            for c in owned:
                col_rows = gi.get_column(c)
                assert np.array_equal(np.isin(rows, col_rows), np.array([c in cols[off[t]:off[t + 1]] for t in range(rows.size)]))
        finally:
            gi.close()


def _fnv1a(data):
    h = 0xcbf29ce484222325
    for b in data:
        h = ((h ^ b) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
    return h


def test_image_with_valid_checksum_but_insane_tables_is_rejected(ggq, tmp_path):
    """The FNV sum only proves the tables are what the writer wrote.  Forge files whose checksum matches but whose
    tables index outside the arrays (child group, leaf rank, slot base, column path): GGQ_ERR_FORMAT, not a crash."""
    import struct
    n, m, d = 4000, 37, 0.05
    gm = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=3, arity=3)
    img = str(tmp_path / "t.ggqimg")
    gm.save_image(img)
    exp = gm.get_rows_csr(np.arange(0, n, 7, dtype=np.uint64))
    gm.close()
    raw = bytearray(open(img, "rb").read())
    hdr = 8 + 8 + 5 * 8 + 3 * 8 + 6 * 4 + 4 * 8   # ImageHeader (ggq_image.hpp)
    # walk the length-prefixed tables to find where each starts and where the checksum sits
    widths = [8, 8, 4, 4, 4, 8, 4, 4, 4, 1, 4, 8, 8, 1, 4]
    pos, starts = hdr, []
    for w in widths:
        cnt = struct.unpack_from("<Q", raw, pos)[0]
        starts.append((pos + 8, cnt, w))
        pos += 8 + cnt * w
    # the host tree: node count, then per node 52 fixed bytes, cols, children, partition groups; then lane_node
    def skip_vec(p, w):
        return p + 8 + struct.unpack_from("<Q", raw, p)[0] * w
    n_nodes = struct.unpack_from("<Q", raw, pos)[0]
    pos += 8
    for _ in range(n_nodes):
        pos += 8 + 7 * 4 + 2 * 8
        pos = skip_vec(skip_vec(pos, 4), 4)
        ng = struct.unpack_from("<Q", raw, pos)[0]
        pos += 8
        for _ in range(ng):
            pos = skip_vec(pos, 4)
    pos = skip_vec(pos, 4)
    sum_at = pos
    assert struct.unpack_from("<Q", raw, sum_at)[0] == _fnv1a(raw[:sum_at])   # the walk matches the writer

    def forged(table, index, value):
        b = bytearray(raw)
        off, cnt, w = starts[table]
        assert index < cnt
        struct.pack_into({1: "<B", 4: "<I", 8: "<Q"}[w], b, off + index * w, value)
        struct.pack_into("<Q", b, sum_at, _fnv1a(b[:sum_at]))
        return bytes(b)

    cases = [forged(0, 1, (1 << 39) | (2 << 40) | (3 << 42)),   # group_hot: slot base far outside the arena
             forged(2, 0, 5_000_000),                            # lane_info: root's child group out of range
             forged(2, starts[2][1] - 4, 0x80000000 | 10_000),   # lane_info: leaf rank beyond the columns
             forged(6, 0, 77_777),                               # leaf_col: column id out of range
             forged(8, 0, 0xFFFFFFF0),                           # col_path: group out of range
             forged(7, 1, 0xFFFFFFF0)]                           # col_path_off: not monotone / beyond the paths
    for bad in cases:
        open(img, "wb").write(bad)
        with pytest.raises(ggq.GgqError) as e:
            ggq.GpuMatrix.load_image(img)
        assert e.value.code == -1
    open(img, "wb").write(bytes(raw))
    gi = ggq.GpuMatrix.load_image(img)
    off, cols = gi.get_rows_csr(np.arange(0, n, 7, dtype=np.uint64))
    assert np.array_equal(off, exp[0]) and np.array_equal(cols, exp[1])
    gi.close()


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "brwt_*.npz"))))
def test_stream_files(ggq, path, tmp_path):
    """ggq_brwt_serialize_file / ggq_brwt_load_file (buffered stdio, no copy of the stream in memory): the file is the
    reference's stream byte for byte, and a matrix loaded from it answers like the reference."""
    g = np.load(path)
    data = g["stream"].tobytes()
    gm = ggq.GpuMatrix.load_brwt(data)
    f = str(tmp_path / "m.brwt")
    try:
        assert gm.serialize_to_file(f) == len(data)
    finally:
        gm.close()
    assert open(f, "rb").read() == data
    gf = ggq.GpuMatrix.load_brwt_file(f)
    try:
        n = int(g["stats"][0])
        off, cols = gf.get_rows_csr(np.arange(n, dtype=np.uint64))
        assert np.array_equal(off, g["row_offsets"]) and np.array_equal(cols, g["row_cols"])
    finally:
        gf.close()
    with pytest.raises(ggq.GgqError):
        ggq.GpuMatrix.load_brwt_file(str(tmp_path / "missing.brwt"))
    open(f, "wb").write(data[: max(len(data) // 2, 1)])   # truncated: load() == false
    with pytest.raises(ggq.GgqError):
        ggq.GpuMatrix.load_brwt_file(f)
