#!/bin/bash
# ncu capture (with warp-stall tables) of the cluster-resident time loop: C1 forward and adjoint kernel, 200 steps each
set -u
mkdir -p gpurun_out
A="bench.py --workload c1 --nt 200 --steps 1 --warmup 3 --no-cpu --no-e2e"
timeout 300 python $A > gpurun_out/c1_short.log 2>&1; echo "c1 short rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:acoustic_small --launch-skip 6 --launch-count 2 \
  -o gpurun_out/small_full -f python $A > gpurun_out/ncu_small.log 2>&1; echo "ncu small rc=$?"
ls -la gpurun_out/small_full.ncu-rep
