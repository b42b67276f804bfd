#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --rows 1e9"
for w in 0 1 0 1; do
  echo "== WIDE_REC=$w"; if [ $w = 1 ]; then export GGQ_WIDE_REC=1; else unset GGQ_WIDE_REC; fi
  timeout 300 $B 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'])"
done
