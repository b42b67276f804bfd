#!/bin/bash
# K1 v3: launch list of the default bench command + one full capture of the descent kernel (1e9-row footprint)
mkdir -p gpurun_out
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 400 $BCMD > gpurun_out/plain_launch_v3.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_v3_c4.csv $BCMD > gpurun_out/ncu_launch_v3.log 2>&1
FCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --rows 1e9"
timeout 400 $FCMD > gpurun_out/plain_full_v3.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:brwt_rows_kernel -s 3 -c 1 -f -o gpurun_out/prof_rows_v3 $FCMD > gpurun_out/ncu_full_v3.log 2>&1
tail -2 gpurun_out/plain_full_v3.log | cut -c1-400
echo done
