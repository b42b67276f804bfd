#!/bin/bash
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_builder_gpu.py -m gpu -q 2>&1 | tail -40 ) 2>&1 | tail -46
