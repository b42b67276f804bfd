#!/bin/bash
# load path at scale: a >= 10 GB stream written from the device image, loaded from the file in a fresh process
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_golden_gpu.py -q -x -k "stream_files or matrix_fixture" 2>&1 | tail -3
F=/tmp/ggq_big.brwt
avail=$(df --output=avail -BG /tmp | tail -1 | tr -dc 0-9)
if [ "$avail" -lt 16 ]; then F=/dev/shm/ggq_big.brwt; fi
memg=$(free -g | awk '/Mem:/{print $7}')
ROWS=7e8
if [ "$memg" -lt 64 ]; then ROWS=1e8; fi
echo "file $F rows $ROWS (avail disk ${avail}G, mem ${memg}G, host threads $(nproc))"
timeout 900 python scripts/load_probe.py big $F --rows $ROWS 2>&1 | tee gpurun_out/r2w_load_big.txt | cut -c1-1200
rm -f $F
