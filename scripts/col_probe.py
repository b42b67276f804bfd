"""get_column at scale: wall time per call and (under ncu) the kernels behind it."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import genome_graph_annotation_b200 as g
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 500_000_000
gm = g.GpuMatrix.synthesize_brwt(n, 2652, 0.003, seed=42, arity=4)
for c in (7, 900, 2651):
    t = time.perf_counter()
    r = gm.get_column(c)
    dt = time.perf_counter() - t
    print("col %d: %d rows in %.1f ms (%.3g rows/s)" % (c, r.size, dt * 1e3, r.size / dt), flush=True)
    assert np.all(np.diff(r.astype(np.int64)) > 0)
gm.close()
