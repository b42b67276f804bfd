#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_v3h_c4.json 2> gpurun_out/bench_v3h_c4.err
tail -3 gpurun_out/bench_v3h_c4.err
python -c "
import json
d = json.loads(open('gpurun_out/bench_v3h_c4.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['other_calls_host_to_host'])"
