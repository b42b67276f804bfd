#!/bin/bash
# the driver's own N=2 command line (bitvector block on rank 0 while the other rank has left)
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 \
  bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_n2_default.json 2> gpurun_out/r2_n2_default.err; echo rc=$?
tail -c 500 gpurun_out/r2_n2_default.err
python - <<'P'
import json
l = json.loads(open("gpurun_out/r2_n2_default.json").read().strip().splitlines()[-1])
print("N=2 value %.4g e2e %.4g cl %.4g c5 %.4g clocks %s / %s bitvector %s" % (l["value"], l["e2e"]["value"], l["count_labels_e2e"]["value"], l["c5"]["value"], l["clocks"], l["c5"]["clocks"], "bitvector" in l))
P
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 \
  bench.py --impl reference --gpus 2 --steps 3 --warmup 1 2>/dev/null | cut -c1-300
