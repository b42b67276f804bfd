#!/bin/bash
# the no-alloc entry point: alone in a fresh process, then at the end of the whole GPU suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -x 2>&1 | tail -5 | tee gpurun_out/r2w_noalloc_alone.txt
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 | tee gpurun_out/r2w_suite2.txt
