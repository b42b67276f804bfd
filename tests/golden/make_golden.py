"""Generates tests/golden/*.npz with the UNMODIFIED reference (oracle/_ref/libggref.so, built from
/root/reference by oracle/Makefile).  Each fixture holds the reference's own serialized stream and the
reference's own answers, so the oracle restatement, the host parser and the CUDA path can be pinned
on boxes where /root/reference does not exist.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import ref_binding as R  # noqa: E402


def matrix_fixture(name, dense, kind="brwt", **kw):
    n, m = dense.shape
    cw = R.pack_columns(dense) if m else np.zeros((0, 1), dtype=np.uint64)
    rm = R.RefMatrix.build_brwt(cw, n, **kw) if kind == "brwt" else R.RefMatrix.build_colmajor(cw, n)
    data = rm.serialize()
    rows = np.arange(n, dtype=np.uint64)
    off, cols, _ = rm.get_rows(rows)
    col_off = [0]
    col_rows = []
    for c in range(m):
        r = rm.get_column(c)
        col_rows.append(r)
        col_off.append(col_off[-1] + r.size)
    stats = [rm.num_rows, rm.num_columns, rm.num_relations]
    if kind == "brwt":
        stats += [rm.num_nodes, rm.total_column_size]
    np.savez_compressed(os.path.join(HERE, name + ".npz"),
                        stream=np.frombuffer(data, dtype=np.uint8), kind=kind,
                        dense=np.packbits(dense, axis=None), shape=np.array(dense.shape),
                        row_offsets=off, row_cols=cols,
                        col_offsets=np.array(col_off, dtype=np.uint64),
                        col_rows=np.concatenate(col_rows) if col_rows else np.zeros(0, np.uint64),
                        stats=np.array(stats, dtype=np.uint64))
    print(name, dense.shape, len(data), "bytes")


def bv_fixture(name, bits):
    bits = np.asarray(bits, dtype=np.uint8)
    n = bits.size
    pad = np.zeros(((n + 63) // 64) * 64, dtype=np.uint8)
    pad[:n] = bits
    words = np.packbits(pad, bitorder="little").view(np.uint64).copy() if n else np.zeros(0, np.uint64)
    out = {"bits": np.packbits(bits), "n": np.array([n])}
    pos = np.arange(n + 3, dtype=np.uint64)
    ones = int(bits.sum())
    ks = np.arange(1, ones + 1, dtype=np.uint64)
    for kind in ("stat", "sd", "rrr"):
        rv = R.RefBitVector.build(kind, words, n)
        out["stream_" + kind] = np.frombuffer(rv.serialize(), dtype=np.uint8)
        out["rank1_" + kind] = rv.rank1(pos)
        out["select1_" + kind] = rv.select1(ks) if ones else np.zeros(0, np.uint64)
        out["get_" + kind] = rv.get(pos[:n]) if n else np.zeros(0, np.uint64)
        out["inverted_" + kind] = np.array([int(rv.is_inverted)])
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, n, ones)


def main():
    rng = np.random.default_rng(20240607)
    i, j = np.meshgrid(np.arange(19), np.arange(19), indexing="ij")
    matrix_fixture("brwt_19x19_checker_arity2", ((i + 2 * j) % 2).astype(np.uint8), arity=2)
    matrix_fixture("brwt_19x19_checker_relaxed", ((i + 2 * j) % 2).astype(np.uint8), arity=2, relax=100)
    matrix_fixture("brwt_empty", np.zeros((0, 0), np.uint8))
    matrix_fixture("brwt_one_column", (rng.random((50, 1)) < 0.5).astype(np.uint8), arity=2)
    d = (rng.random((400, 40)) < 0.08).astype(np.uint8)
    d[7] = 1  # a full row pins the DFS column order
    matrix_fixture("brwt_400x40_greedy_relax7", d, arity=2, greedy=True, relax=7)
    d = (rng.random((300, 65)) < 0.3).astype(np.uint8)
    d[0] = 1
    matrix_fixture("brwt_300x65_arity4", d, arity=4)
    d = (rng.random((200, 30)) < 0.2).astype(np.uint8)
    d[:, 3] = 1
    d[:, 4] = 0
    matrix_fixture("colmajor_200x30", d, kind="colmajor")
    bv_fixture("bv_golden16", [0, 1, 0, 1, 1, 1, 1, 0, 0, 1, 0, 0, 0, 0, 1, 1])
    bv_fixture("bv_zeros10", np.zeros(10, np.uint8))
    bv_fixture("bv_ones10", np.ones(10, np.uint8))
    bv_fixture("bv_2017_sparse", (rng.random(2017) < 0.02).astype(np.uint8))
    bv_fixture("bv_5000_dense", (rng.random(5000) < 0.8).astype(np.uint8))


if __name__ == "__main__":
    main()
