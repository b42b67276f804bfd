#!/bin/bash
# Round 2, call C: compact host call (tests + bench e2e), launch list of the bench command, full ncu capture of the
# coded descent at the 1e9-row footprint.
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) 2>&1 | tail -10
timeout 500 python bench.py --steps 10 --warmup 3 > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err
tail -c 800 gpurun_out/r2c_bench.err
python - <<'P'
import json
try:
    l = json.loads(open("gpurun_out/r2c_bench.json").read().strip().splitlines()[-1])
    r = l["roofline"]
    print("value %.4g" % l["value"], "kernel_ms %.3f" % r["kernel_ms"], "e2e %.4g" % l["e2e"]["value"], l["e2e"]["api"][:24], "u32 %.4g" % l["e2e_u32"]["value"],
          "u64 %.4g" % l["e2e_u64"]["value"], "cl %.4g" % l["count_labels_e2e"]["value"], "cpu", l.get("cpu_baseline", {}).get("value"), l.get("cpu_baseline", {}).get("cores"))
    for b in l.get("bitvector", {}).get("results", []):
        print(b["kind"], b["density"], {k: round(b[k]["gops_per_s"], 1) for k in ("get", "rank1", "select1")}, {k: round(b[k]["roofline_frac"], 2) for k in ("get", "rank1", "select1")}, b.get("identical_to_reference"), b["residency"][:3])
    print("bitvector seconds", l.get("bitvector", {}).get("seconds"))
except Exception as e:
    print("failed:", e)
P
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-bitvector --spot-rows 4096"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2c_launches_c4.csv $BCMD > gpurun_out/r2c_ncu_launch.log 2>&1
FCMD="$BCMD --rows 1e9"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:brwt_rows_kernel -s 9 -c 1 -f -o gpurun_out/r2c_prof_rows $FCMD > gpurun_out/r2c_ncu_full.log 2>&1
tail -3 gpurun_out/r2c_ncu_full.log | cut -c1-300
ls -la gpurun_out/*.ncu-rep
