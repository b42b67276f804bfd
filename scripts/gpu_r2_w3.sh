#!/bin/bash
# C5 on one GPU in the capacity layout WITHOUT leaf codes (the bare image), then the whole GPU suite (regression)
mkdir -p gpurun_out
GGQ_NO_CODES=1 timeout 600 python scripts/capacity_c5.py --ab-rows 0 > gpurun_out/r2w_capacity_nocodes.json 2> gpurun_out/r2w_capacity_nocodes.err; echo rc=$?
tail -c 600 gpurun_out/r2w_capacity_nocodes.err
cat gpurun_out/r2w_capacity_nocodes.json | cut -c1-1500
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -15 | tee gpurun_out/r2w_suite.txt
