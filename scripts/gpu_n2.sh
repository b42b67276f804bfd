#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v3_c4_n2.json 2> gpurun_out/bench_v3_c4_n2.err
tail -c 1500 gpurun_out/bench_v3_c4_n2.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v3_c5_n2.json 2> gpurun_out/bench_v3_c5_n2.err
tail -c 2500 gpurun_out/bench_v3_c5_n2.json
