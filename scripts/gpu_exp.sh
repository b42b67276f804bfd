#!/bin/bash
mkdir -p gpurun_out
for c in 20 21 22; do
GGQ_CHUNK_LOG2=$c timeout 600 python bench.py --no-cpu-baseline --rows 1e9 2>/dev/null | python -c "
import json,sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('chunk', $c, d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], 'e2e', d['e2e']['value'], 'u64', d['e2e_u64']['value'], 'cl', d['count_labels_e2e']['value'])"
done
