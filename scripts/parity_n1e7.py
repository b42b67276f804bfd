#!/usr/bin/env python
"""SURVEY.md 8(d) "C4 / C5 concrete input", second half: the C4 generator at n' rows (default 1e7 x 2652, d = 0.3 %) fed to
the UNMODIFIED reference builder — greedy partitioner + relax(7), the reference's "Multi-BRWT" — on the host cores; the
serialized tree is loaded by the engine and EVERY row is compared with the reference's own get_row (order included).
The device builder (ggq_brwt_build_columns) builds the same tree from the same columns and must agree node by node.
Reports build / load seconds and rows/s of both sides on that tree.

  python scripts/parity_n1e7.py [--rows 1e7] [--cols 2652] [--density 0.003] [--out profiles/r02/parity_n1e7.json]
"""
import argparse
import json
import os
import resource
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=float, default=1e7)
    ap.add_argument("--cols", type=int, default=2652)
    ap.add_argument("--density", type=float, default=0.003)
    ap.add_argument("--relax", type=int, default=7)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    threads = max(1, len(os.sched_getaffinity(0)))
    os.environ["OMP_NUM_THREADS"] = str(threads)
    import genome_graph_annotation_b200 as g
    from oracle import ref_binding as R  # checker
    n, m = int(a.rows), a.cols
    thr = g.density_threshold(a.density)
    rec = {"rows": n, "cols": m, "density": a.density, "host_threads": threads, "builder": "greedy + relax(%d)" % a.relax}
    t0 = time.time()
    cw = R.gen_implicit(n, m, thr, 42)
    rec["generate_s"] = time.time() - t0
    t0 = time.time()
    rm = R.RefMatrix.build_brwt(cw, n, arity=2, greedy=True, relax=a.relax, threads=threads)
    rec["reference_build_s"] = time.time() - t0
    data = rm.serialize()
    rec["stream_bytes"] = len(data)
    rec["reference_nodes"] = int(rm.num_nodes)
    t0 = time.time()
    gm = g.GpuMatrix.load_brwt(data)
    rec["engine_load_s"] = time.time() - t0
    rec["engine_load_peak_rss_gb"] = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1e6
    rec["device_image_gb"] = gm.device_bytes / 1e9
    rec["leaf_codes"] = gm.leaf_codes
    assert (gm.num_rows, gm.num_columns, gm.num_relations, gm.num_nodes) == (n, m, rm.num_relations, rm.num_nodes)
    # every row, in chunks
    step = 1 << 21
    cpu_s = gpu_s = 0.0
    ids = 0
    for lo in range(0, n, step):
        rows = np.arange(lo, min(n, lo + step), dtype=np.uint64)
        eoff, ecols, secs = rm.get_rows(rows, threads=threads)
        cpu_s += secs
        t0 = time.time()
        off, cols = gm.get_rows_csr(rows)
        gpu_s += time.time() - t0
        if not (np.array_equal(off, eoff) and np.array_equal(cols, ecols)):
            raise SystemExit("rows %d.. differ from the reference" % lo)
        ids += int(cols.size)
    rec["all_rows_identical"] = True
    rec["column_ids"] = ids
    rec["reference_rows_per_s_in_order"] = n / cpu_s
    rec["engine_rows_per_s_in_order_host_to_host"] = n / gpu_s
    rnd = np.random.default_rng(42).integers(0, n, 1 << 22).astype(np.uint64)
    eoff, ecols, secs = rm.get_rows(rnd, threads=threads)
    gm.get_rows_csr(rnd)   # warm-up at full size: the pinned staging buffer grows once
    t0 = time.time()
    off, cols = gm.get_rows_csr(rnd)
    rec["engine_rows_per_s_random_host_to_host"] = rnd.size / (time.time() - t0)
    rec["reference_rows_per_s_random"] = rnd.size / secs
    assert np.array_equal(off, eoff) and np.array_equal(cols, ecols)
    for c in (0, m // 2, m - 1):
        assert np.array_equal(gm.get_column(c), rm.get_column(c))
    ref_shape = gm.tree_shape()
    gm.close()
    # the device builder on the same columns
    t0 = time.time()
    gd = g.GpuMatrix.build_brwt(cw, n, greedy=True, relax=a.relax)
    rec["device_build_s"] = time.time() - t0
    dev_shape = gd.tree_shape()
    rec["device_builder_same_tree"] = bool(all(np.array_equal(ref_shape[k], dev_shape[k]) for k in ref_shape))
    off, cols = gd.get_rows_csr(rnd)
    rec["device_builder_same_rows"] = bool(np.array_equal(off, eoff) and np.array_equal(cols, ecols))
    t0 = time.time()
    rec["device_built_matrix_serializes_to_the_reference_stream"] = bool(gd.serialize() == data)
    rec["serialize_from_device_s"] = time.time() - t0
    gd.close()
    # the implicit-generator form (no materialised columns at all)
    t0 = time.time()
    gi = g.GpuMatrix.synthesize_brwt_greedy(n, m, a.density, seed=42, relax=a.relax)
    rec["implicit_device_build_s"] = time.time() - t0
    imp_shape = gi.tree_shape()
    rec["implicit_builder_same_tree"] = bool(all(np.array_equal(ref_shape[k], imp_shape[k]) for k in ref_shape))
    gi.close()
    print(json.dumps(rec))
    if a.out:
        json.dump(rec, open(a.out, "w"), indent=1)
    if not (rec["device_builder_same_tree"] and rec["device_builder_same_rows"] and rec["implicit_builder_same_tree"] and
            rec["device_built_matrix_serializes_to_the_reference_stream"]):
        raise SystemExit("device builder differs from the reference builder")


if __name__ == "__main__":
    main()
