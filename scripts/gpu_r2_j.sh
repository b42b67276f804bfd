#!/bin/bash
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -12 ) 2>&1 | tail -18
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err
tail -c 600 gpurun_out/r2j_bench.err
python - <<'P'
import json
try:
    l = json.loads(open("gpurun_out/r2j_bench.json").read().strip().splitlines()[-1])
    r = l["roofline"]
    print("value %.4g" % l["value"], "ms/step %.3f" % l["ms_per_step"], "kernel_ms %.3f" % r["kernel_ms"], "plain %.3f" % r.get("plain_descent_kernel_ms", -1),
          "e2e %.4g" % l["e2e"]["value"], "u32 %.4g" % l["e2e_u32"]["value"], "cl %.4g" % l["count_labels_e2e"]["value"], "cpu", l.get("cpu_baseline", {}).get("value"))
except Exception as e:
    print("failed:", e)
P
