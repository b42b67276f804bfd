#!/usr/bin/env python
"""Chunk size of the pipelined host calls (GGQ_CHUNK_LOG2) on config C4, through ggq_get_rows_compact: rows/s host to
host for each value.  Experiment, not a bench.   python scripts/chunk_sweep.py --chunk-log2 19,20,21,22"""
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
    ap.add_argument("--chunk-log2", default="19,20,21,22")
    args = ap.parse_args()
    import torch
    import genome_graph_annotation_b200 as g
    n = int(args.rows)
    gm = g.GpuMatrix.synthesize_brwt(n, 2652, 0.003, seed=42, arity=4)
    B = args.batch
    rng = np.random.default_rng(1)
    rows = [torch.from_numpy(rng.integers(0, n, B).astype(np.int64).astype(np.int32, casting="unsafe") if False else
                             rng.integers(0, n, B).astype(np.uint32).view(np.int32)).pin_memory() for _ in range(3)]
    np_rows = [r.numpy().view(np.uint32) for r in rows]
    h_cnt = torch.empty(B, dtype=torch.uint8).pin_memory()
    h_cols = torch.empty(B * 9, dtype=torch.int16).pin_memory()
    np_cnt, np_cols = h_cnt.numpy(), h_cols.numpy().view(np.uint16)
    for cl in args.chunk_log2.split(","):
        os.environ["GGQ_CHUNK_LOG2"] = cl
        for i in range(3):
            gm.get_rows_compact(np_rows[i % 3], np_cnt, np_cols)
        ts = []
        for i in range(6):
            t0 = time.perf_counter()
            gm.get_rows_compact(np_rows[i % 3], np_cnt, np_cols)
            ts.append(time.perf_counter() - t0)
        print("chunk 2^%s: best %.3f ms median %.3f ms -> %.4g rows/s" % (cl, min(ts) * 1e3, float(np.median(ts)) * 1e3, B / float(np.median(ts))), flush=True)
    gm.close()


if __name__ == "__main__":
    main()
