#!/bin/bash
# Round 2, call A: GPU tests, default vs flat emit on C4, dram-bytes capture of K1 on the C4 shape.
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 ) 2>&1 | tail -8
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
timeout 300 $B > gpurun_out/r2a_default.json 2> gpurun_out/r2a_default.err
GGQ_EMIT_FLAT=1 timeout 300 $B > gpurun_out/r2a_flat.json 2> gpurun_out/r2a_flat.err
python - <<'P'
import json
for name in ("default", "flat"):
    try:
        l = json.loads(open(f"gpurun_out/r2a_{name}.json").read().strip().splitlines()[-1])
        print(name, "value %.4g" % l["value"], "kernel_ms %.3f" % l["roofline"]["kernel_ms"], "frac %.3f" % l["roofline"]["frac"],
              "e2e %.4g" % l["e2e"]["value"], "cl %.4g" % l["count_labels_e2e"]["value"], "spot", l.get("spot_check", {}).get("identical"))
    except Exception as e:
        print(name, "failed:", e)
P
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum \
  --clock-control none -k regex:brwt_rows_kernel --launch-skip 2 --launch-count 1 --csv --log-file gpurun_out/r2a_traffic_c4.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/r2a_ncu.log 2>&1
tail -3 gpurun_out/r2a_traffic_c4.csv
