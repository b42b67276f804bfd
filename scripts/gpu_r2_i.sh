#!/bin/bash
# Round 2, call I: chunk sweep of the compact host call, K5 sweep (replaces the stale r01 file) + ncu of the K5 kernels,
# DRAM traffic of the coded K1 on the C4 shape.
mkdir -p gpurun_out
timeout 600 python scripts/chunk_sweep.py 2>&1 | tail -6 | tee gpurun_out/r2i_chunk_sweep.txt
timeout 900 python scripts/bench_bv.py --out gpurun_out/r2i_bench_bv_c3.json > gpurun_out/r2i_bench_bv.log 2>&1; tail -c 300 gpurun_out/r2i_bench_bv.log
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread \
  --clock-control none -k regex:"stat_|bv_query" --csv --log-file gpurun_out/r2i_ncu_k5.csv \
  python scripts/bench_bv.py --no-cpu --densities 0.01,0.5 --log2-queries 24 > gpurun_out/r2i_ncu_k5.log 2>&1
grep -c "stat_\|bv_query" gpurun_out/r2i_ncu_k5.csv
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum \
  --clock-control none -k regex:brwt_rows_kernel --launch-skip 9 --launch-count 1 --csv --log-file gpurun_out/r2i_traffic_c4.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-bitvector --spot-rows 4096 > gpurun_out/r2i_ncu_traffic.log 2>&1
grep brwt_rows gpurun_out/r2i_traffic_c4.csv | awk -F'","' '{print $(NF-2), $NF}'
