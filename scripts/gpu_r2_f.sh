#!/bin/bash
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -12 ) 2>&1 | tail -18
# per-sequence regime: C4-sized matrix is not needed for the plumbing; a 5e8 x 2652 matrix keeps the build short
for T in 8 32; do tests/cpp/bin/batcher_test 500000000 2652 0.003 $T 400 300 2>&1 | tail -2 | tee -a gpurun_out/r2f_batcher.txt; done
tests/cpp/bin/batcher_test 1000000 100 0.01 16 400 100 2>&1 | tail -2 | tee -a gpurun_out/r2f_batcher.txt
