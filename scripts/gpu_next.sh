#!/bin/bash
# First GPU call of the next round: time the experimental flat emit (GGQ_EMIT_FLAT=1) against the default on the
# headline configuration, with the tests first.  ~5 GPU-minutes.
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
timeout 300 $B > gpurun_out/bench_next_default.json 2> gpurun_out/bench_next_default.err
GGQ_EMIT_FLAT=1 timeout 300 $B > gpurun_out/bench_next_flat.json 2> gpurun_out/bench_next_flat.err
python - <<'P'
import json
for name in ("default", "flat"):
    try:
        l = json.loads(open(f"gpurun_out/bench_next_{name}.json").read().strip().splitlines()[-1])
        print(name, "value %.4g" % l["value"], "kernel_ms %.3f" % l["roofline"]["kernel_ms"], "frac %.3f" % l["roofline"]["frac"],
              "e2e %.4g" % l["e2e"]["value"], "spot", l.get("spot_check", {}).get("ok"))
    except Exception as e:
        print(name, "failed:", e)
P
# chunk size of the pipelined host calls after the pipeline rework (2^20 was tuned for the old pipeline)
timeout 200 python scripts/trace_e2e.py --chunk-log2 19,20 2> /dev/null
