#!/bin/bash
# latency form of the direct label path: parity tests, then A/B against the copy form on the per-sequence benchmark
mkdir -p gpurun_out; : > gpurun_out/r2w_batcher.txt
timeout 600 python -m pytest tests/test_synth_gpu.py tests/test_brwt_gpu.py tests/test_adapter_gpu.py tests/test_wide_gpu.py -q -x -k "label or batch or annotator or concurrent or same_answers or layouts or golden" 2>&1 | tail -4
for FORM in 0 1; do
  for T in 8 1 32; do
    echo "GGQ_DIRECT_COPY=$FORM threads=$T" | tee -a gpurun_out/r2w_batcher.txt
    GGQ_DIRECT_COPY=$FORM timeout 300 tests/cpp/bin/batcher_test 500000000 2652 0.003 $T 400 300 2>&1 | tail -2 | tee -a gpurun_out/r2w_batcher.txt
  done
done
echo "C1 shape, 16 threads, batch 100" | tee -a gpurun_out/r2w_batcher.txt
timeout 300 tests/cpp/bin/batcher_test 1000000 100 0.01 16 400 100 2>&1 | tail -2 | tee -a gpurun_out/r2w_batcher.txt
