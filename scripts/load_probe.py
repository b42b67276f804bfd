#!/usr/bin/env python
"""Load path at scale: seconds and peak host RSS of loading a reference-serialized BRWT stream (BRWT::load's format,
BRWT.cpp:87-111) through ggq_brwt_load_buffer, lazy (default: vectors stay RRR-compressed on the host until each is
uploaded) vs eager (GGQ_LOAD_EAGER=1: every node decoded to plain bits first, as in round 1).

  python scripts/load_probe.py make  <file> [--rows 3e7] [--cols 2652] [--density 0.003]   # reference builder, basic arity 4
  python scripts/load_probe.py load  <file>                                                  # one JSON line
  python scripts/load_probe.py both  <file> ...                                              # make, then load lazy and eager in fresh processes
  python scripts/load_probe.py big   <file> [--rows 7e8]   # >= 10 GB: the stream is written by ggq_brwt_serialize_file from a
                                                           # device-synthesized matrix (byte-identical to the reference's own
                                                           # serialization of that matrix: tests/test_builder_gpu.py), then
                                                           # loaded from the file in a fresh process and spot-checked
"""
import argparse
import json
import os
import resource
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def make(a):
    threads = max(1, len(os.sched_getaffinity(0)))
    os.environ["OMP_NUM_THREADS"] = str(threads)
    import genome_graph_annotation_b200 as g
    from oracle import ref_binding as R  # builds the input only
    n = int(a.rows)
    t0 = time.time()
    cw = R.gen_implicit(n, a.cols, g.density_threshold(a.density), 42)
    rm = R.RefMatrix.build_brwt(cw, n, arity=4, threads=threads)
    data = rm.serialize()
    open(a.file, "wb").write(data)
    print(json.dumps({"made": a.file, "rows": n, "cols": a.cols, "stream_bytes": len(data), "nodes": int(rm.num_nodes),
                      "total_column_size_bits": int(rm.total_column_size), "reference_build_s": time.time() - t0}))


def load(a):
    import numpy as np
    import genome_graph_annotation_b200 as g
    data = open(a.file, "rb").read()
    rss0 = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1e6
    t0 = time.time()
    gm = g.GpuMatrix.load_brwt(data)
    secs = time.time() - t0
    rss1 = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1e6
    rows = np.random.default_rng(1).integers(0, gm.num_rows, 1 << 20).astype(np.uint64)
    off, cols = gm.get_rows_csr(rows)
    print(json.dumps({"mode": "eager" if os.environ.get("GGQ_LOAD_EAGER") else "lazy", "stream_gb": len(data) / 1e9,
                      "load_s": secs, "peak_rss_gb_before_load": rss0, "peak_rss_gb_after_load": rss1,
                      "host_gb_added_by_load": rss1 - rss0, "device_image_gb": gm.device_bytes / 1e9,
                      "plain_bits_gb": gm.total_column_size / 8e9, "check_ids": int(cols.size)}))
    gm.close()


def make_big(a):
    import genome_graph_annotation_b200 as g
    n = int(a.rows)
    t0 = time.time()
    gm = g.GpuMatrix.synthesize_brwt(n, a.cols, a.density, seed=42, arity=4)
    t1 = time.time()
    nbytes = gm.serialize_to_file(a.file)
    t2 = time.time()
    print(json.dumps({"made": a.file, "rows": n, "cols": a.cols, "stream_bytes": int(nbytes), "nodes": int(gm.num_nodes),
                      "total_column_size_bits": int(gm.total_column_size), "device_build_s": t1 - t0,
                      "serialize_from_device_s": t2 - t1, "serialize_gb_per_s": nbytes / 1e9 / (t2 - t1),
                      "serialize_peak_rss_gb": resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1e6}))
    gm.close()


def load_big(a):
    import numpy as np
    import genome_graph_annotation_b200 as g
    import bench
    size = os.path.getsize(a.file)
    rss0 = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1e6
    t0 = time.time()
    gm = g.GpuMatrix.load_brwt_file(a.file)
    secs = time.time() - t0
    rss1 = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1e6
    rows = np.random.default_rng(1).integers(0, gm.num_rows, 1 << 16).astype(np.uint64)
    chk = bench.spot_check(gm, rows, a.cols, a.density)
    print(json.dumps({"mode": "lazy, from file", "stream_gb": size / 1e9, "load_s": secs, "load_gb_per_s": size / 1e9 / secs,
                      "peak_rss_gb_before_load": rss0, "peak_rss_gb_after_load": rss1, "host_gb_added_by_load": rss1 - rss0,
                      "device_image_gb": gm.device_bytes / 1e9, "plain_bits_gb": gm.total_column_size / 8e9,
                      "rows": int(gm.num_rows), "nodes": int(gm.num_nodes), "spot_check": chk}))
    gm.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["make", "load", "both", "big", "make_big", "load_big"])
    ap.add_argument("file")
    ap.add_argument("--rows", type=float, default=3e7)
    ap.add_argument("--cols", type=int, default=2652)
    ap.add_argument("--density", type=float, default=0.003)
    a = ap.parse_args()
    if a.what == "make_big":
        make_big(a)
    elif a.what == "load_big":
        load_big(a)
    elif a.what == "big":
        args = [sys.executable, os.path.abspath(__file__)]
        tail = [a.file, "--rows", str(a.rows), "--cols", str(a.cols), "--density", str(a.density)]
        try:
            subprocess.check_call(args + ["make_big"] + tail)
            subprocess.check_call(args + ["load_big"] + tail)
        finally:
            if os.path.exists(a.file):
                os.remove(a.file)
    elif a.what == "make":
        make(a)
    elif a.what == "load":
        load(a)
    else:
        args = [sys.executable, os.path.abspath(__file__)]
        subprocess.check_call(args + ["make", a.file, "--rows", str(a.rows), "--cols", str(a.cols), "--density", str(a.density)])
        subprocess.check_call(args + ["load", a.file])
        subprocess.check_call(args + ["load", a.file], env=dict(os.environ, GGQ_LOAD_EAGER="1"))
        os.remove(a.file)


if __name__ == "__main__":
    main()
