"""CPU suite (no GPU): the C-ABI library loads, exports every symbol include/ggq.h declares, fails loudly
without a device, and its host-side logic (stream parser + flattener) agrees with the golden fixtures."""
import ctypes as C
import glob
import os
import re

import numpy as np
import pytest

from conftest import ROOT

GOLD = os.path.join(ROOT, "tests", "golden")


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ggq.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ggq_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(ggq):
    L = ggq.lib()
    names = declared_symbols()
    assert len(names) >= 40
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    # the ctypes binding covers the whole header too
    unbound = [n for n in names if n not in L._ggq_signatures]
    assert not unbound, unbound


def test_no_cpu_fallback(ggq):
    if ggq.device_count() > 0:
        pytest.skip("a CUDA device is present")
    g = np.load(os.path.join(GOLD, "brwt_400x40_greedy_relax7.npz"))
    with pytest.raises(ggq.GgqError) as e:
        ggq.GpuMatrix.load_brwt(g["stream"].tobytes())
    assert e.value.code == -2  # GGQ_ERR_CUDA: the engine refuses to compute without a device
    with pytest.raises(ggq.GgqError):
        ggq.GpuBitVector.build("stat", np.zeros(2, np.uint64), 100)
    with pytest.raises(ggq.GgqError):
        ggq.GpuMatrix.synthesize_brwt(100, 10, 0.1)


def test_image_loader_rejects_non_images_without_a_device(ggq, tmp_path):
    """ggq_image_load validates the header before it touches the device: a missing file, a serialized BRWT stream
    (not an image) and a short file are GGQ_ERR_FORMAT on any box."""
    g = np.load(os.path.join(GOLD, "brwt_400x40_greedy_relax7.npz"))
    not_image = tmp_path / "m.brwt"
    not_image.write_bytes(g["stream"].tobytes())
    short = tmp_path / "short.ggqimg"
    short.write_bytes(b"GGQIMG01" + b"\0" * 10)
    for path in (str(tmp_path / "missing.ggqimg"), str(not_image), str(short)):
        with pytest.raises(ggq.GgqError) as e:
            ggq.GpuMatrix.load_image(path)
        assert e.value.code == -1, path


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "brwt_*.npz"))))
def test_host_flattener_vs_golden(ggq, path):
    g = np.load(path)
    data = g["stream"].tobytes()
    info = ggq.inspect_brwt(data)
    n, m, rel, nodes, tcs = [int(x) for x in g["stats"]]
    assert info["consumed"] == len(data)
    assert (info["num_rows"], info["num_columns"], info["num_relations"], info["num_nodes"],
            info["total_column_size"]) == (n, m, rel, nodes, tcs)
    # DFS (= get_row) order of the columns: any all-ones row of the fixture returns exactly that order
    off, cols = g["row_offsets"], g["row_cols"]
    full = [r for r in range(n) if off[r + 1] - off[r] == m]
    if full and m:
        r = full[0]
        assert np.array_equal(info["dfs_columns"].astype(np.uint64), cols[off[r]:off[r + 1]])
    assert sorted(info["dfs_columns"].tolist()) == list(range(m))
    # every row of the reference is a subsequence of the DFS order
    pos = {int(c): i for i, c in enumerate(info["dfs_columns"])}
    for r in range(n):
        p = [pos[int(c)] for c in cols[off[r]:off[r + 1]]]
        assert p == sorted(p)


def test_host_parser_rejects_bad_streams(ggq):
    g = np.load(os.path.join(GOLD, "brwt_300x65_arity4.npz"))
    data = g["stream"].tobytes()
    for bad in (b"", data[:10], data[: len(data) // 3], b"\xff" * 64, data[:-1]):
        with pytest.raises(ggq.GgqError) as e:
            ggq.inspect_brwt(bad)
        assert e.value.code == -1
    # 200 zero bytes parse as the empty matrix — the reference's BRWT::load accepts them too (checked live)
    z = ggq.inspect_brwt(b"\x00" * 200)
    assert (z["num_rows"], z["num_columns"], z["num_nodes"]) == (0, 0, 1)
    # trailing bytes after a complete matrix are not consumed (BinaryMatrix::load reads exactly its bytes)
    info = ggq.inspect_brwt(data + b"trailing")
    assert info["consumed"] == len(data)


def test_density_threshold_and_mix64(ggq):
    assert ggq.density_threshold(0.0) == 0
    assert ggq.density_threshold(1.0) == 2**64 - 1
    assert ggq.density_threshold(0.5) == 2**63
    # splitmix64 finaliser known answers
    x = ggq.mix64(np.array([0, 1], dtype=np.uint64))
    assert int(x[0]) == 0 and int(x[1]) == 0x5692161D100B05E5
