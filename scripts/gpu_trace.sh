#!/bin/bash
set -u
mkdir -p gpurun_out
FWI_B200_TRACE=$PWD/gpurun_out/trace.txt FWI_B200_PAIR=0 timeout 300 python bench.py --one-batch --nt 24 --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/trace_run.log 2>&1
echo "rc=$?"; ls gpurun_out/trace.txt.*
