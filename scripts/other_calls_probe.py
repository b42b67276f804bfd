"""K2 / K3 / K4 under ncu: get (cells), has_column, get_column on a BRWT and get_row / get_column on a column-major
matrix, so that `ncu --metrics ... -k regex:"brwt_get|lane_|colmajor"` has something to capture.  Prints wall times."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import genome_graph_annotation_b200 as g

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
gm = g.GpuMatrix.synthesize_brwt(n, 2652, 0.003, seed=42, arity=4)
rng = np.random.default_rng(1)
rows = rng.integers(0, n, 1 << 22).astype(np.uint64)
cols = rng.integers(0, 2652, 1 << 22).astype(np.uint64)
gm.get_cells(rows, cols)
t = time.perf_counter(); gm.get_cells(rows, cols); print("get: %.3g cells/s" % (rows.size / (time.perf_counter() - t)), flush=True)
for c in (7, 2651):
    t = time.perf_counter(); r = gm.get_column(c); print("get_column %d: %d rows, %.2f ms" % (c, r.size, (time.perf_counter() - t) * 1e3), flush=True)
gm.close()
from oracle import ref_binding as R  # builds a column-major stream only
nc, mc = 2_000_000, 200
cw = R.gen_norepl(nc, mc, 0.01, seed=42)
cm = R.RefMatrix.build_colmajor(cw, nc)
gc = g.GpuMatrix.load_colmajor(cm.serialize())
rr = rng.integers(0, nc, 1 << 18).astype(np.uint64)
gc.get_rows_csr(rr)
t = time.perf_counter(); gc.get_rows_csr(rr); print("colmajor get_rows: %.3g rows/s" % (rr.size / (time.perf_counter() - t)), flush=True)
t = time.perf_counter(); r = gc.get_column(3); print("colmajor get_column: %d rows, %.2f ms" % (r.size, (time.perf_counter() - t) * 1e3), flush=True)
gc.close()
