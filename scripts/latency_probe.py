#!/usr/bin/env python
"""Per-call latency of the scalar / per-sequence calls (through ctypes, one host thread): get_row (ggq_get_rows, n = 1),
get_rows of 100 rows, get_labels of 300 indices.  GGQ_DIRECT_COPY=1 selects the copy / memset / synchronise form."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import genome_graph_annotation_b200 as g  # noqa: E402


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
    gm = g.GpuMatrix.synthesize_brwt(n, 2652, 0.003, seed=42, arity=4)
    rng = np.random.default_rng(1)
    out = {"form": "copy" if os.environ.get("GGQ_DIRECT_COPY") else "latency", "rows": n}
    for name, k, fn in (("get_row_us", 1, lambda r: gm.get_rows_csr(r)), ("get_rows_100_us", 100, lambda r: gm.get_rows_csr(r)),
                        ("get_labels_300_us", 300, lambda r: gm.get_labels(r, 0.5))):
        qs = [rng.integers(0, n, k).astype(np.uint64) for _ in range(3000)]
        for q in qs[:500]:
            fn(q)
        t0 = time.perf_counter()
        for q in qs[500:]:
            fn(q)
        out[name] = (time.perf_counter() - t0) / 2500 * 1e6
    # the answers of the two forms must be the same: checksum over a fixed set of calls
    chk = 0
    for s in range(200):
        r = np.random.default_rng(100 + s).integers(0, n, 1 + s % 97).astype(np.uint64)
        off, cols = gm.get_rows_csr(r)
        c, k = gm.get_labels(r, 0.3)
        chk = (chk * 1000003 + int(off.sum()) + int(cols.sum()) * 7 + int(c.sum()) * 13 + int(k.sum()) * 17) % (1 << 61)
    out["checksum"] = chk
    print(json.dumps(out))
    gm.close()


if __name__ == "__main__":
    main()
