"""Exploration: time the device builder and the descent kernel at growing row counts (C4 shape)."""
import sys
import time

import numpy as np
import torch

import genome_graph_annotation_b200 as g

m = int(sys.argv[2]) if len(sys.argv) > 2 else 2652
d = float(sys.argv[3]) if len(sys.argv) > 3 else 0.003
B = 1 << 24
for n in [int(float(x)) for x in sys.argv[1].split(",")]:
    torch.cuda.synchronize()
    t0 = time.time()
    gm = g.GpuMatrix.synthesize_brwt(n, m, d, seed=42, arity=4)
    torch.cuda.synchronize()
    t1 = time.time()
    print("n=%.3g build %.2fs nodes=%d total_column_size=%.4g bits/row=%.2f device_bytes=%.3f GB relations/row=%.3f" % (
        n, t1 - t0, gm.num_nodes, gm.total_column_size, gm.total_column_size / n, gm.device_bytes / 1e9,
        gm.num_relations / n), flush=True)
    ws = gm.workspace()
    gen = torch.Generator(device="cuda")
    gen.manual_seed(42)
    rows = torch.randint(0, n, (B,), device="cuda", dtype=torch.int64, generator=gen)
    st = torch.cuda.ExternalStream(ws.stream)
    for it in range(4):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            e0.record()
            off, cols, total = ws.get_rows_device(rows.data_ptr(), B, 4)
            e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        probes = gm.last_probe_count
        print("  iter %d: %.3f ms  %.3g rows/s  ids=%d probes/row=%.2f  probe-bytes(64B) %.1f GB/s" % (
            it, ms, B / ms * 1e3, total, probes / B, probes * 64 / ms / 1e6), flush=True)
    ws.close()
    gm.close()
    del rows
    torch.cuda.empty_cache()
