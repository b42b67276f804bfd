#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_c4_n2.json 2> gpurun_out/bench_c4_n2.err
echo "exit $?" >> gpurun_out/bench_c4_n2.err
timeout 600 python -m pytest tests/test_shard_gpu.py -m gpu -x -q -k nccl > gpurun_out/pytest_nccl.log 2>&1
echo done
