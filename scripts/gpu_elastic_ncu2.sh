#!/bin/bash
# Round-2 ncu evidence for the elastic step kernels (C3, one shot group so that launches are in step order):
# launch list + one --set full capture (with warp-stall tables) of two forward and two adjoint steps.
set -u
mkdir -p gpurun_out
export FWI_B200_EL_GROUPS=1
A="bench.py --workload c3 --nt 40 --steps 1 --warmup 3 --no-cpu --no-e2e"
timeout 300 python $A > gpurun_out/c3_short.log 2>&1; echo "c3 short rc=$?"; tail -1 gpurun_out/c3_short.log | cut -c1-200
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:elastic_ --launch-skip 720 -c 240 --csv \
  --log-file gpurun_out/launches_c3.csv python $A > gpurun_out/ncu_launches_c3.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:elastic_f --launch-skip 300 --launch-count 2 \
  -o gpurun_out/el_fwd_full -f python $A > gpurun_out/ncu_el_fwd.log 2>&1; echo "ncu el fwd rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:elastic_b --launch-skip 300 --launch-count 2 \
  -o gpurun_out/el_adj_full -f python $A > gpurun_out/ncu_el_adj.log 2>&1; echo "ncu el adj rc=$?"
ls -la gpurun_out/*.ncu-rep
