#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi2.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 scripts/shard_nccl_check.py > gpurun_out/nccl_check.log 2>&1
echo "exit $?" >> gpurun_out/nccl_check.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --workload c5 --steps 5 --warmup 3 > gpurun_out/bench_c5_n2.json 2> gpurun_out/bench_c5_n2.err
echo "exit $?" >> gpurun_out/bench_c5_n2.err
echo done
