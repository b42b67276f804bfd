#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python scripts/bench_c12.py > gpurun_out/r2o_bench_c12.json 2> gpurun_out/r2o_bench_c12.err; tail -c 300 gpurun_out/r2o_bench_c12.err
python - <<'P'
import json
for l in open("gpurun_out/r2o_bench_c12.json"):
    try:
        j = json.loads(l)
        print(j["config"][:40], "all %.3g/%.3g" % (j["get_rows_all"]["gpu_rows_per_s_host_to_host"], j["get_rows_all"]["cpu_reference_rows_per_s"]),
              {k: "%.3g" % v for k, v in j["get_rows_random"]["gpu_rows_per_s_by_batch_size"].items()}, "cpu rnd %.3g" % j["get_rows_random"]["cpu_reference_rows_per_s"],
              "col %.3f/%.2f ms" % (j["get_column"]["gpu_ms_per_column"], j["get_column"]["cpu_reference_ms_per_column"]), "row %.1f us" % j["get_row_single_call_us"],
              "build ref %.1f dev %.1f" % (j["reference_build_s"], j["device_build_s"]), j["leaf_codes"]["nodes"])
    except Exception as e:
        print("bad line", e)
P
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread \
  --clock-control none -k regex:"brwt_get|lane_|colmajor" --csv --log-file gpurun_out/r2o_ncu_k234.csv python scripts/other_calls_probe.py 1e9 > gpurun_out/r2o_ncu_k234.log 2>&1
tail -6 gpurun_out/r2o_ncu_k234.log; grep -c "brwt_get\|lane_\|colmajor" gpurun_out/r2o_ncu_k234.csv
