#!/bin/bash
mkdir -p gpurun_out
for T in 32 32 8 1; do tests/cpp/bin/batcher_test 500000000 2652 0.003 $T 400 300 2>&1 | tail -3 | tee -a gpurun_out/r2m_batcher.txt; done
tests/cpp/bin/batcher_test 1000000 100 0.01 16 400 100 2>&1 | tail -2 | tee -a gpurun_out/r2m_batcher.txt
timeout 600 python -m pytest tests/test_synth_gpu.py tests/test_adapter_gpu.py -m gpu -q 2>&1 | tail -3
