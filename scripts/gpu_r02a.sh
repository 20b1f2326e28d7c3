#!/bin/bash
# Round-2 first GPU pass: parity tests (incl. BASELINE-config and drop-in tests), smoke, bench line (32-shot survey
# slice), shots-per-launch sweep, ncu launch list + full captures (stall tables) of the two acoustic step kernels.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/host.txt; free -g >> gpurun_out/host.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?" | tee -a gpurun_out/bench.log
tail -2 gpurun_out/bench.log
for s in 3 4 5 6 7; do
  timeout 300 python bench.py --shots $s --nt 300 --steps 2 --no-cpu --no-e2e > gpurun_out/bench_shots$s.log 2>&1
  echo "shots $s rc=$?"; tail -1 gpurun_out/bench_shots$s.log | cut -c1-1200
done
CMD="python bench.py --nt 40 --steps 1 --warmup 3 --no-cpu --no-e2e --one-batch"
$CMD > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:acoustic_step \
    --launch-skip 20 --launch-count 1 -o gpurun_out/fwd_full -f $CMD > gpurun_out/ncu_fwd.log 2>&1; echo "ncu fwd rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:acoustic_step \
    --launch-skip 60 --launch-count 1 -o gpurun_out/adj_full -f $CMD > gpurun_out/ncu_adj.log 2>&1; echo "ncu adj rc=$?"
ls -la gpurun_out
