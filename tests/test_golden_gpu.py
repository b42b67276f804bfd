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
    for c in range(m):
        assert np.array_equal(gm.get_column(c), crows[coff[c]:coff[c + 1]])
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
