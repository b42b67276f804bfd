#!/bin/bash
# trimmed final run (K1 itself unchanged since the last full ncu capture): bench line, reference arm, launch list
mkdir -p gpurun_out
timeout 400 python bench.py > gpurun_out/bench_final2_c4.json 2> gpurun_out/bench_final2_c4.err
tail -c 400 gpurun_out/bench_final2_c4.json
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_final2_ref.json 2> gpurun_out/bench_final2_ref.err
tail -c 300 gpurun_out/bench_final2_ref.json
timeout 140 python __graft_entry__.py smoke > gpurun_out/smoke_final2.log 2>&1; tail -1 gpurun_out/smoke_final2.log
echo done
