#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 250 python scripts/trace_e2e.py --chunk-log2 20 > gpurun_out/trace5_e2e.log 2> gpurun_out/trace5_e2e.err; cat gpurun_out/trace5_e2e.log
timeout 200 python - <<'P'
import numpy as np, time, torch
import genome_graph_annotation_b200 as g
n=int(3.7e9)
gm = g.GpuMatrix.synthesize_brwt(n, 2652, 0.003, seed=42, arity=4)
rng=np.random.default_rng(2)
rows=torch.from_numpy(rng.integers(0,n,1<<24).astype(np.int64)).pin_memory().numpy().view(np.uint64)
for i in range(2): c=gm.count_labels(rows)
best=1e9
for i in range(4):
    t=time.perf_counter(); c2=gm.count_labels(rows); best=min(best,time.perf_counter()-t)
assert np.array_equal(c,c2)
print("count_labels 2^24: %.2f ms %.3e rows/s sum %d"%(best*1e3,(1<<24)/best,int(c.sum())))
P
