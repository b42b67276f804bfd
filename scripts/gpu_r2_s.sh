#!/bin/bash
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_builder_gpu.py -m gpu -q -x -k "implicit" 2>&1 | tail -25 ) 2>&1 | tail -30
timeout 900 python bench.py --tree greedy_relax7 --steps 10 --warmup 3 --no-cpu-baseline --no-bitvector > gpurun_out/r2s_bench_greedy.json 2> gpurun_out/r2s_bench_greedy.err
tail -c 600 gpurun_out/r2s_bench_greedy.err
python - <<'P'
import json
try:
    l = json.loads(open("gpurun_out/r2s_bench_greedy.json").read().strip().splitlines()[-1])
    r = l["roofline"]
    print("greedy tree: value %.4g" % l["value"], "kernel_ms %.3f" % r["kernel_ms"], "plain %.3f" % r.get("plain_descent_kernel_ms", -1), "touches/row %.1f" % r["node_touches_per_row"],
          "probes %.2f code %.2f esc %.2f" % (r["probes_per_row"], r["leaf_code_probes_per_row"], r["leaf_code_escapes_per_row"]), "e2e %.4g" % l["e2e"]["value"],
          "img %.1f GB" % l["device_image_gb"], l["leaf_codes"], "build %.0f s" % l["build_seconds"], l["spot_check"]["identical"])
except Exception as e:
    print("failed:", e)
P
