#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -12
for b in 1048576 4194304; do
timeout 600 python bench.py --no-cpu-baseline --no-e2e --rows 1e9 --batch $b --steps 20 2>/dev/null | python -c "
import json,sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('batch', $b, d['value'], d['ms_per_step'], d['roofline']['kernel_ms'])"
done
