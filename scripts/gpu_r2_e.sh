#!/bin/bash
mkdir -p gpurun_out
oracle/_ref/annotator_test 2>&1 | tail -20
timeout 300 python -m pytest tests/test_synth_gpu.py -m gpu -x -q -k "compact or labels_batch or batch" 2>&1 | tail -8
FCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-bitvector --spot-rows 4096 --rows 1e9"
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,launch__occupancy_limit_shared_mem,launch__occupancy_limit_registers,sm__warps_active.avg.pct_of_peak_sustained_active,launch__shared_mem_per_block_dynamic,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:brwt_rows_kernel -s 8 -c 2 --csv --log-file gpurun_out/r2e_k1_metrics.csv $FCMD > gpurun_out/r2e_ncu.log 2>&1
python - <<'P'
import csv
for r in csv.reader(open("gpurun_out/r2e_k1_metrics.csv")):
    if len(r) > 10 and "brwt_rows" in r[4]: print(r[0], r[-3], r[-2], r[-1])
P
