#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v3_c4_n$N.json 2> gpurun_out/bench_v3_c4_n$N.err
tail -c 1200 gpurun_out/bench_v3_c4_n$N.json; tail -3 gpurun_out/bench_v3_c4_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v3_c5_n${N}_peer.json 2> gpurun_out/bench_v3_c5_n${N}_peer.err
tail -c 1500 gpurun_out/bench_v3_c5_n${N}_peer.json; tail -3 gpurun_out/bench_v3_c5_n${N}_peer.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus $N --workload c5 --no-peer --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v3_c5_n${N}_nccl.json 2> gpurun_out/bench_v3_c5_n${N}_nccl.err
tail -c 1500 gpurun_out/bench_v3_c5_n${N}_nccl.json; tail -3 gpurun_out/bench_v3_c5_n${N}_nccl.err
