#!/bin/bash
# Round 2, call D: summary-word emit + 6 CTAs/SM again, annotator test, compact-call failure message.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_synth_gpu.py -m gpu -x -q -k "compact" 2>&1 | tail -30
( time timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -15 ) 2>&1 | tail -22
timeout 500 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-bitvector > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err
tail -c 800 gpurun_out/r2d_bench.err
python - <<'P'
import json
try:
    l = json.loads(open("gpurun_out/r2d_bench.json").read().strip().splitlines()[-1])
    r = l["roofline"]
    print("value %.4g" % l["value"], "ms/step %.3f" % l["ms_per_step"], "kernel_ms %.3f" % r["kernel_ms"], "plain %.3f" % r.get("plain_descent_kernel_ms", -1),
          "e2e %.4g" % l["e2e"]["value"], "u32 %.4g" % l["e2e_u32"]["value"], "cl %.4g" % l["count_labels_e2e"]["value"], l["other_calls_host_to_host"]["small_calls"])
except Exception as e:
    print("failed:", e)
P
