#!/bin/bash
# C5 unsharded on one B200 in the capacity layout + the layout's price on the C4 column geometry
mkdir -p gpurun_out
timeout 1200 python scripts/capacity_c5.py > gpurun_out/r2w_capacity.json 2> gpurun_out/r2w_capacity.err; echo rc=$?
tail -c 1500 gpurun_out/r2w_capacity.err
python - <<'P'
import json
r = json.loads(open("gpurun_out/r2w_capacity.json").read().strip().splitlines()[-1])
print(json.dumps(r, indent=1)[:6000])
P
