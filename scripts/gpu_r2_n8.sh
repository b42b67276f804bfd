#!/bin/bash
# Round 2, 8 GPUs: 8-rank NCCL / peer / slice parity, then the default bench at N=8 (C4 replicas + the c5 block).
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \
  scripts/shard_nccl_check.py > gpurun_out/r2_n8_nccl_check.log 2>&1; echo "nccl check rc=$?"; grep -c "slice peer" gpurun_out/r2_n8_nccl_check.log; tail -2 gpurun_out/r2_n8_nccl_check.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 \
  bench.py --gpus 8 --steps 5 --warmup 3 --no-bitvector > gpurun_out/r2_n8_bench.json 2> gpurun_out/r2_n8_bench.err
tail -c 800 gpurun_out/r2_n8_bench.err
python - <<'P'
import json
try:
    l = json.loads(open("gpurun_out/r2_n8_bench.json").read().strip().splitlines()[-1])
    print("C4 N=8 value %.4g e2e %.4g (%s) u32 %.4g cl %.4g clocks %s" % (l["value"], l["e2e"]["value"], l["e2e"]["api"][:20], l["e2e_u32"]["value"], l["count_labels_e2e"]["value"], l["clocks"]))
    c = l["c5"]
    print("C5 value %.4g ms %.3f kernel %.3f exchange %s clocks %s spot %s img %.1f GB" % (c["value"], c["ms_per_step"], c["roofline"]["kernel_ms_max_over_ranks"], c["exchange"], c["clocks"], c["spot_check"]["identical"], c["device_image_gb_per_rank_max"]))
except Exception as e:
    print("failed:", e)
P
