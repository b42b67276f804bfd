#!/bin/bash
# Round 2, call B: leaf codes — new parity tests, C4 bench (codes on; the line carries the plain-descent time), gather
# ceiling between 64 and 128 GiB footprints.
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) 2>&1 | tail -10
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-bitvector > gpurun_out/r2b_codes.json 2> gpurun_out/r2b_codes.err
tail -c 600 gpurun_out/r2b_codes.err
python - <<'P'
import json
try:
    l = json.loads(open("gpurun_out/r2b_codes.json").read().strip().splitlines()[-1])
    r = l["roofline"]
    print("value %.4g" % l["value"], "kernel_ms %.3f" % r["kernel_ms"], "plain %.3f" % r.get("plain_descent_kernel_ms", -1), "frac %.3f" % r["frac"],
          "probes/row %.2f" % r["probes_per_row"], "code probes %.2f esc %.3f" % (r["leaf_code_probes_per_row"], r["leaf_code_escapes_per_row"]),
          "e2e %.4g" % l["e2e"]["value"], "cl %.4g" % l["count_labels_e2e"]["value"], "img %.1f GB" % l["device_image_gb"], l["leaf_codes"], "build %.0f s" % l["build_seconds"], l["spot_check"])
except Exception as e:
    print("failed:", e)
P
timeout 300 scripts/mb_gather 124 0 64 80 96 104 112 120 124 > gpurun_out/r2b_mb_footprints.txt 2>&1
cat gpurun_out/r2b_mb_footprints.txt
