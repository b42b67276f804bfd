#!/usr/bin/env python
"""Device timeline of one pipelined ggq_get_rows_into_u32 call on config C4 (GGQ_TRACE=1, see ggq_engine.cu):
where the host-to-host step spends its time.  Debug aid, not a bench.

  python scripts/trace_e2e.py [--rows 3.7e9] [--chunk-log2 20]
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=float, default=3.7e9)
    ap.add_argument("--batch", type=int, default=1 << 24)
    ap.add_argument("--chunk-log2", default="20")
    args = ap.parse_args()
    import torch
    import genome_graph_annotation_b200 as g
    n = int(args.rows)
    gm = g.GpuMatrix.synthesize_brwt(n, 2652, 0.003, seed=42, arity=4)
    B = args.batch
    rng = np.random.default_rng(1)
    rows = [torch.from_numpy(rng.integers(0, n, B).astype(np.int64)).pin_memory() for _ in range(2)]
    np_rows = [r.numpy().view(np.uint64) for r in rows]
    h_off = torch.empty(B + 1, dtype=torch.int64).pin_memory()
    h_cols = torch.empty(B * 9, dtype=torch.int32).pin_memory()
    np_off, np_cols = h_off.numpy().view(np.uint64), h_cols.numpy().view(np.uint32)
    for cl in args.chunk_log2.split(","):
        os.environ["GGQ_CHUNK_LOG2"] = cl
        for i in range(3):
            gm.get_rows_into(np_rows[i % 2], np_off, np_cols)
        best = 1e9
        for i in range(4):
            t = time.perf_counter()
            gm.get_rows_into(np_rows[i % 2], np_off, np_cols)
            best = min(best, time.perf_counter() - t)
        print(f"chunk 2^{cl}: {best * 1e3:.2f} ms per call, {B / best:.3e} rows/s", flush=True)
        os.environ["GGQ_TRACE"] = "1"
        sys.stderr.write(f"==== chunk 2^{cl}\n"); sys.stderr.flush()
        gm.get_rows_into(np_rows[0], np_off, np_cols)
        del os.environ["GGQ_TRACE"]
    gm.close()


if __name__ == "__main__":
    main()
