#!/usr/bin/env python
"""Configs C1 and C2 of BASELINE.json (SURVEY.md 8d "C1 / C2 concrete input") on one B200, with the unmodified
reference timed beside it on the box's host cores.

  C1  experiments brwt on simulated norepl 1,000,000 x 100 (d = 0.01, DataGenerator seed 42): "brwt_arity_2" and
      "brwt_greedy_relax_7"; get_rows over all rows in order + the reference's sampled mode
  C2  Multi-BRWT (greedy + relax 7) on simulated uniform_rows 1,000,000 x 1000 (2e5 unique rows x 5, d = 0.01):
      batched get_rows over all rows and over 1e6 random ids at batch sizes 1e2 / 1e4 / 1e6, get_column for 10
      random columns and for all 1000

The matrices are generated and built by the reference's own code (oracle/_ref: data_generation.cpp +
BRWT_builders.cpp), serialized with the reference's serialize() and loaded through ggq_brwt_load_buffer.  Every GPU
answer that is timed is also compared with the reference's (all rows, all columns): a mismatch aborts.  One JSON
line per configuration.

  python scripts/bench_c12.py [--rows 1000000] [--configs c1a,c1b,c2]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, reps=3):
    fn()
    best = 1e30
    for _ in range(reps):
        t = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t)
    return best


def run(name, n, m, d, gen, build_kw, args, g, R):
    threads = R.max_threads()
    t0 = time.time()
    cw = gen()
    rm = R.RefMatrix.build_brwt(cw, n, threads=threads, **build_kw)
    build_s = time.time() - t0
    data = rm.serialize()
    t0 = time.time()
    gm = g.GpuMatrix.load_brwt(data)
    load_s = time.time() - t0
    line = {"config": name, "rows": n, "cols": m, "density": d, "builder": build_kw, "reference_build_s": build_s,
            "stream_bytes": len(data), "gpu_load_s": load_s, "device_image_bytes": int(gm.device_bytes),
            "num_nodes": int(gm.num_nodes), "avg_arity": float(rm.avg_arity),
            "node_touches_per_row": float(gm.total_column_size) / n, "cpu_threads": threads,
            "leaf_codes": gm.leaf_codes}
    # the device builder on the same columns: same tree as the reference builder's, and how long it takes
    t0 = time.time()
    gd = g.GpuMatrix.build_brwt(cw, n, greedy=build_kw.get("greedy", False), arity=build_kw.get("arity", 2),
                                relax=build_kw.get("relax", 0))
    line["device_build_s"] = time.time() - t0
    sa, sb = gd.tree_shape(), gm.tree_shape()
    assert all(np.array_equal(sa[k], sb[k]) for k in sa), "device builder: tree differs from the reference builder's"
    line["device_builder_same_tree"] = True
    gd.close()
    # ---- get_rows over all rows in order (parity on every row) ---------------------------------------------
    all_rows = np.arange(n, dtype=np.uint64)
    eoff, ecols, cpu_s = rm.get_rows(all_rows, threads=threads)
    off, cols = gm.get_rows_csr(all_rows)
    assert np.array_equal(off, eoff) and np.array_equal(cols, ecols), "get_rows(all) differs from the reference"
    h_off = np.empty(n + 1, np.uint64)
    h_cols = np.empty(max(int(eoff[-1]), 1), np.uint32)
    gpu_s = timed(lambda: gm.get_rows_into(all_rows, h_off, h_cols))
    cpu_s = min(cpu_s, rm.time_get_rows(all_rows, threads)[0])
    line["get_rows_all"] = {"gpu_rows_per_s_host_to_host": n / gpu_s, "cpu_reference_rows_per_s": n / cpu_s,
                            "column_ids": int(eoff[-1])}
    # ---- random ids, batch sizes -----------------------------------------------------------------------------
    rnd = R.gen_random_ints(n, 0, n, seed=42)
    roff, rcols, _ = rm.get_rows(rnd, threads=threads)
    cpu_rnd_s = rm.time_get_rows(rnd, threads)[0]
    per_batch = {}
    h_rcols = np.empty(max(int(roff[-1]), 1), np.uint32)
    for bs in (100, 10_000, n):
        nb = max(1, min(n // bs, 200))
        def go():
            for b in range(nb):
                if bs >= 10_000:   # caller-provided buffers, as a real bulk caller would use
                    gm.get_rows_into(rnd[b * bs:(b + 1) * bs], h_off[:bs + 1], h_rcols)
                else:
                    gm.get_rows_csr(rnd[b * bs:(b + 1) * bs])
        o, c = gm.get_rows_csr(rnd[:bs])
        assert np.array_equal(o, roff[:bs + 1]) and np.array_equal(c, rcols[:int(roff[bs])]), "random batch differs"
        per_batch[str(bs)] = nb * bs / timed(go, reps=2)
    line["get_rows_random"] = {"gpu_rows_per_s_by_batch_size": per_batch, "cpu_reference_rows_per_s": n / cpu_rnd_s}
    # the reference's sampled mode: 1000 rows, one call each (experiments/main.cpp:78-93)
    sample = R.gen_random_ints(1000, 0, n, seed=42)
    t = time.perf_counter()
    for r in sample[:200]:
        gm.get_row(int(r))
    line["get_row_single_call_us"] = (time.perf_counter() - t) / 200 * 1e6
    # ---- get_column ------------------------------------------------------------------------------------------
    pick = [int(c) for c in R.gen_random_ints(10, 0, m, seed=42)]
    t = time.perf_counter()
    ref_cols = {c: rm.get_column(c) for c in pick}
    cpu_col_s = (time.perf_counter() - t) / len(pick)
    for c in pick:
        assert np.array_equal(gm.get_column(c), ref_cols[c]), "get_column differs"
    gpu_col_s = timed(lambda: [gm.get_column(c) for c in pick]) / len(pick)
    t = time.perf_counter()
    total = 0
    for c in range(m):
        total += gm.get_column(c).size
    all_cols_s = time.perf_counter() - t
    assert total == int(gm.num_relations)
    line["get_column"] = {"gpu_ms_per_column": gpu_col_s * 1e3, "cpu_reference_ms_per_column": cpu_col_s * 1e3,
                          "gpu_all_columns_s": all_cols_s, "rows_out_per_s": total / all_cols_s}
    line["parity"] = "all rows, 1e6 random rows, 10 + all-columns count: identical to the unmodified reference"
    gm.close()
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--configs", default="c1a,c1b,c2")
    args = ap.parse_args()
    import genome_graph_annotation_b200 as g
    from oracle import ref_binding as R   # checker + CPU baseline
    if g.device_count() < 1:
        raise SystemExit("bench_c12: no CUDA device (the engine has no CPU fallback)")
    if not R.available():
        raise SystemExit("bench_c12: oracle/_ref is not built")
    n = args.rows
    want = set(args.configs.split(","))
    if "c1a" in want:
        run("C1 norepl 1e6 x 100 d=0.01, brwt_arity_2", n, 100, 0.01, lambda: R.gen_norepl(n, 100, 0.01, seed=42),
            dict(arity=2), args, g, R)
    if "c1b" in want:
        run("C1 norepl 1e6 x 100 d=0.01, brwt_greedy_relax_7", n, 100, 0.01, lambda: R.gen_norepl(n, 100, 0.01, seed=42),
            dict(arity=2, greedy=True, relax=7), args, g, R)
    if "c2" in want:
        run("C2 uniform_rows 1e6 x 1000 d=0.01 (2e5 unique x 5), Multi-BRWT greedy + relax 7", n, 1000, 0.01,
            lambda: R.gen_uniform_rows(n, 1000, 0.01, n // 5, seed=42), dict(arity=2, greedy=True, relax=7), args, g, R)


if __name__ == "__main__":
    main()
