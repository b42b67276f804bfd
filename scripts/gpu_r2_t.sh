#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_bitvector_gpu.py tests/test_golden_gpu.py -m gpu -q -x 2>&1 | tail -4
timeout 600 python scripts/bench_bv.py --kinds rrr --no-cpu --densities 0.001,0.01,0.1 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try:
        r=json.loads(l); print('inline', r['kind'], r['density'], {o: round(v['ops_per_s']/1e9,1) for o,v in r['ops'].items()}, 'bits/bit %.3f' % r['bits_per_bit'])
    except Exception: pass
"
GGQ_RRR_NO_INLINE=1 timeout 600 python scripts/bench_bv.py --kinds rrr --no-cpu --densities 0.001,0.01 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try:
        r=json.loads(l); print('split ', r['kind'], r['density'], {o: round(v['ops_per_s']/1e9,1) for o,v in r['ops'].items()}, 'bits/bit %.3f' % r['bits_per_bit'])
    except Exception: pass
"
