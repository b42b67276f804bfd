#!/usr/bin/env python
"""Capacity layout at full scale on ONE B200.

  1. C5 (RefSeq-shaped 1e9 x 3173, 3.8 %, arity 4): 182.8 GB in the default layout — more than one GPU holds, which is why
     the benchmark shards it over 8 — synthesized here in the capacity layout (32-byte slots of 224 positions), UNSHARDED, on
     one device; device-resident get_rows over batches of 2^22 uniformly random rows, and 2^14 rows of the answers checked
     against the direct evaluation of the implicit matrix.
  2. The price of the layout where both fit: the C4 column geometry at --ab-rows rows in both layouts, same batches.

One JSON object on stdout.  Single GPU, no torch.distributed."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402  (direct_rows, SEED)


def measure(g, torch, gm, n, B, K, W, m, d, spot):
    ws = gm.workspace()
    stream = torch.cuda.ExternalStream(ws.stream, device=torch.device("cuda", 0))
    gen = torch.Generator(device="cuda")
    batches = []
    for b in range(min(K + W, 4)):
        gen.manual_seed(bench.SEED + b)
        batches.append(torch.randint(0, n, (B,), device="cuda", dtype=torch.int64, generator=gen))
    for i in range(W):
        ws.get_rows_device(batches[i % len(batches)].data_ptr(), B, 8)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kern, ids = [], 0
    with torch.cuda.stream(stream):
        e0.record()
    for i in range(K):
        _, _, total = ws.get_rows_device(batches[(W + i) % len(batches)].data_ptr(), B, 8)
        kern.append(ws.stat("kernel_ms"))
        ids += total
    with torch.cuda.stream(stream):
        e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    out = {"rows_per_s": B / (ms / 1e3), "ms_per_step": ms, "descent_kernel_ms": float(np.mean(kern)),
           "column_ids_per_row": ids / (B * K), "batch_rows": B, "steps": K, "warmup": W}
    if spot:
        rows = batches[0][:spot].cpu().numpy().view(np.uint64)
        out["spot_check"] = bench.spot_check(gm, rows, m, d)
    ws.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=float, default=1e9)
    ap.add_argument("--cols", type=int, default=3173)
    ap.add_argument("--density", type=float, default=0.038)
    ap.add_argument("--batch", type=int, default=1 << 22)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--spot-rows", type=int, default=1 << 14)
    ap.add_argument("--ab-rows", type=float, default=1e9, help="rows of the C4-geometry A/B (0: skip)")
    ap.add_argument("--no-c5", action="store_true")
    a = ap.parse_args()
    import torch
    import genome_graph_annotation_b200 as g
    torch.cuda.set_device(0)
    free0, total = torch.cuda.mem_get_info()
    res = {"device_memory_gb": total / 1e9}
    if not a.no_c5:
        n, m, d = int(a.rows), a.cols, a.density
        g.set_default_option("wide_slots", 1)
        t0 = time.time()
        gm = g.GpuMatrix.synthesize_brwt(n, m, d, seed=bench.SEED, arity=4)
        torch.cuda.synchronize()
        blk = {"workload": "C5 RefSeq-shaped synthetic BRWT %.3g x %d density %g, arity-4 basic partitioner, UNSHARDED on one "
                           "B200 in the capacity layout (32-byte slots of 224 positions)" % (n, m, d),
               "slot_bits": int(gm.slot_bits), "build_seconds": time.time() - t0, "device_image_gb": gm.device_bytes / 1e9,
               "default_layout_image_gb": 182.8, "nodes": int(gm.num_nodes), "leaf_codes": gm.leaf_codes,
               "free_after_build_gb": torch.cuda.mem_get_info()[0] / 1e9}
        blk.update(measure(g, torch, gm, n, a.batch, a.steps, a.warmup, m, d, a.spot_rows))
        gm.close()
        res["c5_one_gpu"] = blk
    if a.ab_rows:
        n, m, d = int(a.ab_rows), 2652, 0.003
        ab = {"workload": "C4 column geometry (2652 columns, 0.3 %%, arity 4) at %.3g rows, batches of 2^24 random rows" % n}
        for name, wide in (("default_layout", 0), ("capacity_layout", 1)):
            g.set_default_option("wide_slots", wide)
            gm = g.GpuMatrix.synthesize_brwt(n, m, d, seed=bench.SEED, arity=4)
            r = {"slot_bits": int(gm.slot_bits), "device_image_gb": gm.device_bytes / 1e9, "leaf_codes": gm.leaf_codes}
            r.update(measure(g, torch, gm, n, 1 << 24, a.steps, a.warmup, m, d, 1 << 14))
            gm.set_option("leaf_codes", 0)
            r["without_leaf_codes"] = measure(g, torch, gm, n, 1 << 24, a.steps, a.warmup, m, d, 0)
            gm.close()
            ab[name] = r
        g.set_default_option("wide_slots", 0)
        res["c4_geometry_ab"] = ab
    print(json.dumps(res))


if __name__ == "__main__":
    main()
