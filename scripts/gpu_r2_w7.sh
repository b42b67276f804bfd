#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_brwt_gpu.py tests/test_golden_gpu.py tests/test_adapter_gpu.py tests/test_synth_gpu.py -q -x 2>&1 | tail -3
for FORM in "" 1; do GGQ_DIRECT_COPY=$FORM timeout 300 python scripts/latency_probe.py 2>&1 | tail -1; done | tee gpurun_out/r2w_latency.txt
