#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python scripts/bench_c12.py > gpurun_out/bench_c12.json 2> gpurun_out/bench_c12.err
tail -c 3000 gpurun_out/bench_c12.json; tail -5 gpurun_out/bench_c12.err
