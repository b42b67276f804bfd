#!/usr/bin/env python
"""Load path at scale: seconds and peak host RSS of loading a reference-serialized BRWT stream (BRWT::load's format,
BRWT.cpp:87-111) through ggq_brwt_load_buffer, lazy (default: vectors stay RRR-compressed on the host until each is
uploaded) vs eager (GGQ_LOAD_EAGER=1: every node decoded to plain bits first, as in round 1).

  python scripts/load_probe.py make  <file> [--rows 3e7] [--cols 2652] [--density 0.003]   # reference builder, basic arity 4
  python scripts/load_probe.py load  <file>                                                  # one JSON line
  python scripts/load_probe.py both  <file> ...                                              # make, then load lazy and eager in fresh processes
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["make", "load", "both"])
    ap.add_argument("file")
    ap.add_argument("--rows", type=float, default=3e7)
    ap.add_argument("--cols", type=int, default=2652)
    ap.add_argument("--density", type=float, default=0.003)
    a = ap.parse_args()
    if a.what == "make":
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
