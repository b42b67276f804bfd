#!/bin/bash
# Round 2, 2 GPUs: NCCL / peer / slice parity (scripts/shard_nccl_check.py via the pytest), then the default bench at N=2
# (C4 replicas + the c5 block).
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_shard_gpu.py -m gpu -q -x 2>&1 | tail -15
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus 2 --steps 5 --warmup 3 --no-bitvector > gpurun_out/r2_n2_bench.json 2> gpurun_out/r2_n2_bench.err
tail -c 1500 gpurun_out/r2_n2_bench.err
python - <<'P'
import json
try:
    l = json.loads(open("gpurun_out/r2_n2_bench.json").read().strip().splitlines()[-1])
    print("C4 N=2 value %.4g e2e %.4g cl %.4g" % (l["value"], l["e2e"]["value"], l["count_labels_e2e"]["value"]))
    c = l["c5"]
    print("C5 value %.4g ms %.3f kernel %.3f exchange %s clocks %s spot %s img %.1f GB" % (c["value"], c["ms_per_step"], c["roofline"]["kernel_ms_max_over_ranks"], c["exchange"], c["clocks"], c["spot_check"], c["device_image_gb_per_rank_max"]))
except Exception as e:
    print("failed:", e)
P
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --gpus 2 --workload c5 --c5-full --steps 5 --warmup 2 > gpurun_out/r2_n2_c5_full.json 2>> gpurun_out/r2_n2_bench.err
python - <<'P'
import json
try:
    c = json.loads(open("gpurun_out/r2_n2_c5_full.json").read().strip().splitlines()[-1])
    print("C5 full value %.4g ms %.3f kernel %.3f exchange %s" % (c["value"], c["ms_per_step"], c["roofline"]["kernel_ms_max_over_ranks"], c["exchange"]))
except Exception as e:
    print("failed:", e)
P
