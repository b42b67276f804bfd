#!/usr/bin/env python
"""Config C3 of BASELINE.json: rank1 / select1 / get micro-benchmark on 2^32-bit
bit_vector_{stat,sd,rrr} at densities 0.1 % .. 50 %, generated on the device with the portable generator
(bit i set iff mix64(seed*0x9E3779B97F4A7C15 + i) < floor(d*2^64)).

Per (kind, density, op): 2^26 uniformly random queries resident in HBM, CUDA events on the launch stream,
>= 3 warm-ups, best-of-5 and median; ops/s and algorithmic GB/s (SURVEY 8d bytes per op) against the measured
HBM peak.  Parity + CPU baseline: the same generator at n' = 2^28 bits feeds the UNMODIFIED reference
(oracle/_ref: bit_vector_stat / bit_vector_sd / bit_vector_rrr<63>); 2^20 sampled answers must be identical and
the reference is timed on all host threads.  One JSON line per (kind, density); a summary JSON at the end.

  python scripts/bench_bv.py [--log2-bits 32] [--log2-queries 26] [--kinds stat,sd,rrr] [--densities 0.001,0.01,0.1,0.5]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

ALG_BYTES = {  # 8 in + 8 out (1 out for get) + sectors touched
    "stat": {"get": 8 + 1 + 32, "rank1": 16 + 32, "select1": 16 + 64},
    "sd": {"get": 8 + 1 + 96, "rank1": 16 + 96, "select1": 16 + 96},
    "rrr": {"get": 8 + 1 + 64, "rank1": 16 + 64, "select1": 16 + 96},
}
SEED = 42


def host_bits(g, n_bits, thr, seed):
    """The generator on the host, word by word (numpy), for the CPU reference."""
    words = np.zeros(n_bits // 64, dtype=np.uint64)
    base = np.uint64((seed * g.SEED_MUL) % 2**64)
    chunk = 1 << 22
    for w0 in range(0, words.size, chunk // 64):
        i = np.arange(w0 * 64, min(n_bits, w0 * 64 + chunk), dtype=np.uint64)
        with np.errstate(over="ignore"):
            bits = g.mix64(base + i) < np.uint64(thr)
        words[w0:w0 + bits.size // 64] = np.packbits(bits.reshape(-1, 64), axis=1, bitorder="little").view(np.uint64).ravel()
    return words


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def measure(kinds, densities, log2_bits=32, log2_queries=26, log2_cpu_bits=28, cpu=True, emit=None):
    """One record per (kind, density).  GPU: 2^log2_queries random queries on a 2^log2_bits-bit vector, CUDA events on
    the launch stream, 3 warm-ups, median of 5.  CPU + parity: the unmodified reference on the 2^log2_cpu_bits-bit
    prefix of the same generator, all host threads, 2^20 sampled answers compared."""
    import torch
    import genome_graph_annotation_b200 as g
    if not torch.cuda.is_available():
        raise RuntimeError("no CUDA device; the engine has no CPU fallback")
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    R = None
    threads = host_threads()
    if cpu:
        os.environ["OMP_NUM_THREADS"] = str(threads)   # torchrun exports 1; set before libgomp loads
        from oracle import ref_binding as R  # checker / baseline only
        if not R.available():
            R = None
    n, Q, nc = 1 << log2_bits, 1 << log2_queries, 1 << log2_cpu_bits
    l2_bytes = 126e6
    gen = torch.Generator(device="cuda")
    results = []
    for d in densities:
        thr = g.density_threshold(d)
        cw = host_bits(g, nc, thr, SEED) if R is not None else None
        for kind in kinds:
            t0 = time.time()
            v = g.GpuBitVector.synthesize(kind, n, d, seed=SEED)
            torch.cuda.synchronize()
            build_s = time.time() - t0
            ones = int(v.num_set_bits)
            rec = {"config": "C3", "kind": kind, "density": d, "n_bits": n, "ones": ones, "queries": Q,
                   "device_bytes": int(v.device_bytes), "bits_per_bit": 8.0 * v.device_bytes / n, "build_s": build_s,
                   "inverted": bool(v.is_inverted),
                   "residency": "L2-resident structure (smaller than the 126 MB L2): not an HBM figure"
                                if v.device_bytes < l2_bytes else "HBM (structure larger than L2)",
                   "ops": {}}
            gen.manual_seed(SEED)
            pos = torch.randint(0, n, (Q,), device="cuda", dtype=torch.int64, generator=gen)
            rk = torch.randint(1, max(ones, 1) + 1, (Q,), device="cuda", dtype=torch.int64, generator=gen)
            out64 = torch.empty(Q, device="cuda", dtype=torch.int64)
            out8 = torch.empty(Q, device="cuda", dtype=torch.uint8)
            st = torch.cuda.current_stream().cuda_stream
            for op in ("get", "rank1", "select1"):
                if op == "select1" and ones == 0:
                    continue
                fn = {"get": lambda: v.get_device(pos.data_ptr(), Q, out8.data_ptr(), st),
                      "rank1": lambda: v.rank1_device(pos.data_ptr(), Q, out64.data_ptr(), st),
                      "select1": lambda: v.select1_device(rk.data_ptr(), Q, out64.data_ptr(), st)}[op]
                for _ in range(3):
                    fn()
                ts = []
                for _ in range(5):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    torch.cuda.synchronize()
                    e0.record(); fn(); e1.record()
                    torch.cuda.synchronize()
                    ts.append(e0.elapsed_time(e1))
                ms = float(np.median(ts))
                gbs = ALG_BYTES[kind][op] * Q / (ms / 1e3) / 1e9
                rec["ops"][op] = {"ms": ms, "ms_best": float(min(ts)), "ops_per_s": Q / (ms / 1e3),
                                  "algorithmic_bytes_per_op": ALG_BYTES[kind][op],
                                  "algorithmic_gb_s": gbs, "frac_of_measured_hbm_peak": gbs / peak}
            v.close()
            # parity + CPU baseline on the n' prefix (same generator => same bits)
            if R is not None:
                vs = g.GpuBitVector.synthesize(kind, nc, d, seed=SEED)
                rv = R.RefBitVector.build(kind, cw, nc)
                rng = np.random.default_rng(SEED)
                qp = rng.integers(0, nc, 1 << 20).astype(np.uint64)
                so = int(vs.num_set_bits)
                qr = rng.integers(1, max(so, 1) + 1, 1 << 20).astype(np.uint64)
                par = {}
                for op, q in (("get", qp), ("rank1", qp), ("select1", qr)):
                    if op == "select1" and so == 0:
                        continue
                    got = getattr(vs, op)(q)
                    exp, secs = rv.timed(op, q, threads)
                    par[op] = bool(np.array_equal(np.asarray(got).astype(np.uint64), np.asarray(exp).astype(np.uint64)))
                    rec["ops"][op]["cpu_reference_ops_per_s"] = q.size / max(secs, 1e-9)
                rec["parity_vs_reference_2^20_samples"] = par
                rec["cpu"] = {"threads": threads, "n_bits": nc, "kind": "reference (oracle/_ref, unmodified sdsl-backed " + kind + ")"}
                vs.close()
                del rv
            if emit is not None:
                emit(rec)
            results.append(rec)
    return peak, results


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2-bits", type=int, default=32)
    ap.add_argument("--log2-queries", type=int, default=26)
    ap.add_argument("--log2-cpu-bits", type=int, default=28)
    ap.add_argument("--kinds", default="stat,sd,rrr")
    ap.add_argument("--densities", default="0.001,0.01,0.1,0.5")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    peak, results = measure(a.kinds.split(","), [float(x) for x in a.densities.split(",")], a.log2_bits, a.log2_queries,
                            a.log2_cpu_bits, cpu=not a.no_cpu, emit=lambda rec: print(json.dumps(rec), flush=True))
    if a.out:
        json.dump({"peak_hbm_gbs": peak, "results": results}, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
