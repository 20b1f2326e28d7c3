#!/bin/bash
# ncu evidence for the elastic kernels (C3) + small-grid (C1) bench.  gpurun --timeout 1200 -- 'bash scripts/gpu_elastic_ncu.sh'
set -u
mkdir -p gpurun_out
timeout 400 python bench.py --workload c3 --no-cpu > gpurun_out/bench_c3.log 2>&1; echo "c3 rc=$?"; tail -1 gpurun_out/bench_c3.log | cut -c1-300
A="bench.py --workload c3 --nt 40 --steps 1 --warmup 3 --no-cpu --no-e2e"
timeout 300 python $A > gpurun_out/c3_short.log 2>&1; echo "c3 short rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:elastic_ --launch-skip 960 -c 320 --csv \
  --log-file gpurun_out/launches_c3.csv python $A > gpurun_out/ncu_launches_c3.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:elastic_ --launch-skip 1000 --launch-count 4 \
  -o gpurun_out/el_fwd_full -f python $A > gpurun_out/ncu_el_fwd.log 2>&1; echo "ncu el fwd rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:elastic_ --launch-skip 1160 --launch-count 4 \
  -o gpurun_out/el_adj_full -f python $A > gpurun_out/ncu_el_adj.log 2>&1; echo "ncu el adj rc=$?"
ls -la gpurun_out/*.ncu-rep
