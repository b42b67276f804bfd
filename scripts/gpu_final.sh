#!/bin/bash
# final build of the round: GPU tests, the default bench line, the reference arm, the ncu launch list of the same
# command, one full capture of the descent kernel (1e9-row footprint: the 3.7e9-row capture would replay 80 GB ~40x)
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_final.log 2>&1; tail -1 gpurun_out/smoke_final.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_final.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_final.log
tail -3 gpurun_out/pytest_final.log
timeout 900 python bench.py > gpurun_out/bench_final_c4.json 2> gpurun_out/bench_final_c4.err
tail -c 600 gpurun_out/bench_final_c4.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_final_ref.json 2> gpurun_out/bench_final_ref.err
tail -c 600 gpurun_out/bench_final_ref.json
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_final_c4.csv $BCMD > gpurun_out/ncu_launch_final.log 2>&1
FCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --rows 1e9"
timeout 400 $FCMD > gpurun_out/plain_full_final.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:brwt_rows_kernel -s 3 -c 1 -f -o gpurun_out/prof_rows_final $FCMD > gpurun_out/ncu_full_final.log 2>&1
echo done
