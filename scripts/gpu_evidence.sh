#!/bin/bash
# Round-2 evidence pass: full GPU test suite, smoke, default bench line, C1/C3/C4/C5/f3 lines, ncu launch list + full
# captures of the acoustic step kernels (forward MODE 1 of a 3-shot group; adjoint MODE 0 + MODE 3 launches).
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1800 python -m pytest tests -m gpu -q --durations=10 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | tail -10
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?" | tee -a gpurun_out/bench.log
tail -1 gpurun_out/bench.log | cut -c1-300
for wl in c1 c3 c4 c5 f3; do
  timeout 900 python bench.py --workload $wl --steps 2 > gpurun_out/bench_$wl.log 2>&1; echo "$wl rc=$?"; tail -1 gpurun_out/bench_$wl.log | cut -c1-200
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.log 2>&1; tail -1 gpurun_out/bench_reference.log | cut -c1-200
if [ "${NCU:-1}" = "1" ]; then
CMD="python bench.py --nt 40 --steps 1 --warmup 3 --no-cpu --no-e2e --one-batch"
$CMD > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv \
    --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:acoustic_step \
    --launch-skip 20 --launch-count 1 -o gpurun_out/fwd_full -f $CMD > gpurun_out/ncu_fwd.log 2>&1; echo "ncu fwd rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:acoustic_step \
    --launch-skip 90 --launch-count 4 -o gpurun_out/adj_full -f $CMD > gpurun_out/ncu_adj.log 2>&1; echo "ncu adj rc=$?"
fi
ls gpurun_out | head -50
