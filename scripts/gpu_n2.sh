#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29510 scripts/shard_nccl_check.py > gpurun_out/nccl_check_peer_n2.log 2>&1
tail -8 gpurun_out/nccl_check_peer_n2.log | cut -c1-300
timeout 600 python -m pytest tests/test_shard_gpu.py -m gpu -x -q 2>&1 | tail -3
