#!/bin/bash
mkdir -p gpurun_out
cd scripts
timeout 200 ./mb_gather 128 > ../gpurun_out/mb_sweep.txt 2>&1
echo done
