#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-bitvector --spot-rows 4096"
timeout 300 $B > gpurun_out/r2k_default.json 2> gpurun_out/r2k_default.err
GGQ_L2_PERSIST=1 timeout 300 $B > gpurun_out/r2k_l2persist.json 2> gpurun_out/r2k_l2persist.err
python - <<'P'
import json
for name in ("default", "l2persist"):
    try:
        l = json.loads(open(f"gpurun_out/r2k_{name}.json").read().strip().splitlines()[-1])
        print(name, "value %.4g" % l["value"], "kernel_ms %.3f" % l["roofline"]["kernel_ms"], "clocks", l["clocks"])
    except Exception as e:
        print(name, "failed:", e, open(f"gpurun_out/r2k_{name}.err").read()[-400:])
P
