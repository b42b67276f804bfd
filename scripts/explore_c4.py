"""Exploration: descent kernel time on the C4 shape under different engine options (one build, many configs)."""
import sys
import time

import numpy as np
import torch

import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import genome_graph_annotation_b200 as g

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else int(3.7e9)
m, d, B = 2652, 0.003, 1 << 24
t0 = time.time()
gm = g.GpuMatrix.synthesize_brwt(n, m, d, seed=42, arity=4)
torch.cuda.synchronize()
print("n=%.3g build %.1fs device_bytes=%.2f GB" % (n, time.time() - t0, gm.device_bytes / 1e9), flush=True)
gen = torch.Generator(device="cuda")
gen.manual_seed(42)
rows = torch.randint(0, n, (B,), device="cuda", dtype=torch.int64, generator=gen)
configs = [(sb, rb) for sb in (0, 8, 12, 16, 20, 64) for rb in (8, 16)]
if len(sys.argv) > 2:
    configs = [tuple(int(x) for x in c.split(":")) for c in sys.argv[2].split(",")]
for sb, rb in configs:
    gm.set_option("sort_bits", sb)
    gm.set_option("rows_per_warp", rb)
    ws = gm.workspace()
    st = torch.cuda.ExternalStream(ws.stream)
    best, bestk = 1e9, 1e9
    for it in range(4):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            e0.record()
            off, cols, total = ws.get_rows_device(rows.data_ptr(), B, 8)
            e1.record()
        torch.cuda.synchronize()
        if it:
            best = min(best, e0.elapsed_time(e1)); bestk = min(bestk, ws.stat("kernel_ms"))
    print("sort_bits=%2d rb=%2d: step %.3f ms  descent %.3f ms  %.3g rows/s  probes/row %.2f ids %d" % (
        sb, rb, best, bestk, B / best * 1e3, ws.stat("probes") / B, total), flush=True)
    ws.close()
gm.close()
