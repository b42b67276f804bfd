#!/bin/bash
mkdir -p gpurun_out
for T in 32 32 1 8; do tests/cpp/bin/batcher_test 500000000 2652 0.003 $T 400 300 2>&1 | tail -4 | tee -a gpurun_out/r2l_batcher.txt; done
tests/cpp/bin/batcher_test 1000000 100 0.01 16 400 100 2>&1 | tail -3 | tee -a gpurun_out/r2l_batcher.txt
( time timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -12 ) 2>&1 | tail -16
