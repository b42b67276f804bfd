#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_v3f_c4.json 2> gpurun_out/bench_v3f_c4.err
python -c "
import json
d = json.loads(open('gpurun_out/bench_v3f_c4.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], 'e2e', d['e2e']['value'], 'u64', d['e2e_u64']['value'], 'cl', d['count_labels_e2e']['value'])"
