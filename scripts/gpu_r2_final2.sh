#!/bin/bash
# closing check of round 2: whole GPU suite, smoke(), the driver's N=1 bench line and its reference arm
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 | tee gpurun_out/r2z_suite.txt
timeout 300 python -c "import __graft_entry__ as e; e.smoke(); print('SMOKE OK')" 2>&1 | tail -2 | tee gpurun_out/r2z_smoke.txt
timeout 900 python bench.py > gpurun_out/r2z_bench.json 2> gpurun_out/r2z_bench.err; echo bench rc=$?
python - <<'P'
import json
l = json.loads(open("gpurun_out/r2z_bench.json").read().strip().splitlines()[-1])
o = l.get("other_calls_host_to_host") or {}
print("value %.4g e2e %.4g frac %.3f clocks %s spot %s" % (l["value"], l["e2e"]["value"], l["roofline"]["frac"], l["clocks"], l["spot_check"]["identical"]))
print({k: o[k] for k in o if "us" in k or "labels" in k or "row" in k})
P
