#!/usr/bin/env python
"""Quick parity + timing check of the experimental EMIT_FLAT descent instantiation (GGQ_EMIT_FLAT=1) against the
default emit on one GPU: identical CSR on a few shapes, then kernel_ms of both on a mid-size matrix."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import genome_graph_annotation_b200 as g


def build(flat, *a, **k):
    if flat:
        os.environ["GGQ_EMIT_FLAT"] = "1"
    else:
        os.environ.pop("GGQ_EMIT_FLAT", None)
    return g.GpuMatrix.synthesize_brwt(*a, **k)


ok = True
for n, m, d, arity in [(20_000, 400, 0.02, 4), (50_000, 2652, 0.003, 4), (9_000, 36, 0.4, 2), (3_000, 1, 0.5, 4)]:
    rows = np.random.default_rng(9).integers(0, n, 40_000).astype(np.uint64)
    rows[::501] = np.uint64(n + 3)
    a = build(False, n, m, d, seed=6, arity=arity)
    base = a.get_rows_csr(rows); small = a.get_rows_csr(rows[:37]); a.close()
    b = build(True, n, m, d, seed=6, arity=arity)
    got = b.get_rows_csr(rows); gsmall = b.get_rows_csr(rows[:37]); b.close()
    same = all(np.array_equal(x, y) for x, y in zip(base + small, got + gsmall))
    ok &= same
    print("shape", n, m, d, arity, "identical" if same else "DIFFERENT", flush=True)
if ok and len(sys.argv) > 1:
    n = int(float(sys.argv[1]))
    rows = np.random.default_rng(1).integers(0, n, 1 << 22).astype(np.uint64)
    off = np.empty(rows.size + 1, np.uint64); cols = np.empty(rows.size * 10, np.uint32)
    for flat in (False, True):
        gm = build(flat, n, 2652, 0.003, seed=42, arity=4)
        ws = gm.workspace()
        best = 1e9
        for i in range(4):
            t = time.perf_counter(); gm.get_rows_into(rows[:1 << 20], off[:(1 << 20) + 1], cols); best = min(best, time.perf_counter() - t)
        print("flat" if flat else "default", "2^20 rows host to host: %.3f ms" % (best * 1e3), flush=True)
        gm.close()
print("flat emit parity:", "ok" if ok else "FAILED")
