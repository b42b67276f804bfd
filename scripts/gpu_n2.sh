#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29510 scripts/shard_nccl_check.py > gpurun_out/nccl_check_peer_n2.log 2>&1
tail -25 gpurun_out/nccl_check_peer_n2.log | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v3_c5_n2_peer.json 2> gpurun_out/bench_v3_c5_n2_peer.err
tail -c 1800 gpurun_out/bench_v3_c5_n2_peer.json; tail -5 gpurun_out/bench_v3_c5_n2_peer.err
