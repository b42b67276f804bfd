#!/bin/bash
# GPU call: parity tests, bench (ours + reference arm), ncu launch list + one full capture of the descent kernel.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
nproc >> gpurun_out/smi.txt
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
echo "bench exit $?" >> gpurun_out/bench_n1.err
timeout 400 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 400 $BCMD > gpurun_out/plain_launch.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $BCMD > gpurun_out/ncu_launch.log 2>&1
FCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --rows 1e9"
timeout 400 $FCMD > gpurun_out/plain_full.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:brwt_rows_kernel -s 3 -c 2 -f -o gpurun_out/prof_rows $FCMD > gpurun_out/ncu_full.log 2>&1
echo done
