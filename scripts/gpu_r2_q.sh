#!/bin/bash
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_golden_gpu.py tests/test_adapter_gpu.py tests/test_shard_gpu.py -m gpu -q -x 2>&1 | tail -25 ) 2>&1 | tail -30
