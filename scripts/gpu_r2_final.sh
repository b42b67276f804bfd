#!/bin/bash
# Round 2, closing run on one GPU: whole -m gpu suite, smoke(), the default bench line + reference arm, ncu launch list.
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -q 2>&1 | tail -8 ) 2>&1 | tail -14
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_final_bench.json 2> gpurun_out/r2_final_bench.err; tail -c 400 gpurun_out/r2_final_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r2_final_ref.json 2> gpurun_out/r2_final_ref.err; tail -c 300 gpurun_out/r2_final_ref.err
python - <<'P'
import json
try:
    l = json.loads(open("gpurun_out/r2_final_bench.json").read().strip().splitlines()[-1])
    r = l["roofline"]
    print("value %.4g" % l["value"], "ms/step %.3f" % l["ms_per_step"], "kernel_ms %.3f" % r["kernel_ms"], "frac %.3f" % r["frac"], "traffic", r["traffic"],
          "e2e %.4g" % l["e2e"]["value"], "cl %.4g" % l["count_labels_e2e"]["value"], "cpu %.4g x%d" % (l["cpu_baseline"]["value"], l["cpu_baseline"]["cores"]), "clocks", l["clocks"], "launches", l["gpu_launches"])
    ref = json.loads(open("gpurun_out/r2_final_ref.json").read().strip().splitlines()[-1])
    print("reference arm %.4g rows/s on %d cores; same config: %s" % (ref["value"], ref["cpu_baseline"]["cores"], ref["config"] == l["config"]))
except Exception as e:
    print("failed:", e)
P
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_final_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-bitvector --spot-rows 4096 > gpurun_out/r2_final_ncu.log 2>&1
python - <<'P'
import csv, collections
agg = collections.OrderedDict()
for r in csv.reader(open("gpurun_out/r2_final_launches.csv")):
    if len(r) > 10 and r[-3] == "gpu__time_duration.sum":
        k = r[4].split("(")[0][-48:]
        a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += float(r[-1]) / 1e6
for k, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
    print("%-50s %5d launches %9.3f ms" % (k, n, ms))
P
