#!/bin/bash
# capacity layout: its own tests + the C++ adapter / annotator programs in both layouts
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_wide_gpu.py tests/test_adapter_gpu.py -q -x 2>&1 | tail -30 | tee gpurun_out/r2w_tests.txt
