#!/bin/bash
# same box, back to back: the library of the earlier closing run (1ddfd2c) and the current one, compact e2e
mkdir -p gpurun_out
show() { python -c "
import json,sys
l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[2], 'value %.4g e2e %.4g u32 %.4g cl %.4g' % (l['value'], l['e2e']['value'], l['e2e_u32']['value'], l['count_labels_e2e']['value']))" $1 $2; }
(cd gpurun_ab/old && timeout 400 python bench.py --no-bitvector --no-cpu-baseline --spot-rows 65536 > ../../gpurun_out/ab_old.json 2>/dev/null); show gpurun_out/ab_old.json old
timeout 400 python bench.py --no-bitvector --no-cpu-baseline --spot-rows 65536 > gpurun_out/ab_new.json 2>/dev/null; show gpurun_out/ab_new.json new
