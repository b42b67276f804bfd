#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python scripts/parity_n1e7.py --out gpurun_out/r2h_parity_n1e7.json 2>&1 | tail -5
