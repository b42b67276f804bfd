#!/bin/bash
# K1 v3: parity tests first, then the C4 bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_v3.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_v3.log
tail -5 gpurun_out/pytest_v3.log
timeout 600 python bench.py > gpurun_out/bench_v3_c4.json 2> gpurun_out/bench_v3_c4.err
tail -c 3000 gpurun_out/bench_v3_c4.json
