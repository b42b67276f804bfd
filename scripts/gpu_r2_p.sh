#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/chunk_sweep.py --chunk-log2 21 2>&1 | tail -1 | tee gpurun_out/r2p_ramp.txt
GGQ_NO_RAMP=1 timeout 600 python scripts/chunk_sweep.py --chunk-log2 20,21,22 2>&1 | tail -3 | tee -a gpurun_out/r2p_ramp.txt
timeout 900 python scripts/bench_c12.py --configs c2 > gpurun_out/r2p_bench_c2.json 2> gpurun_out/r2p_bench_c2.err; tail -c 300 gpurun_out/r2p_bench_c2.err
python - <<'P'
import json
for l in open("gpurun_out/r2p_bench_c2.json"):
    j = json.loads(l)
    print(j["config"][:40], {k: "%.3g" % v for k, v in j["get_rows_random"]["gpu_rows_per_s_by_batch_size"].items()})
P
