#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench line, ncu launch list + full captures of the two
# step kernels.  Run as: gpurun --timeout 1500 -- 'bash scripts/gpu_check.sh'
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?" | tee -a gpurun_out/bench.log
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; tail -2 gpurun_out/bench.log
if [ "${NCU:-1}" = "1" ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv python bench.py --nt 40 --steps 1 --warmup 3 --no-cpu --no-e2e \
    > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:acoustic_step \
    --launch-skip 20 --launch-count 1 -o gpurun_out/fwd_full -f python bench.py --nt 40 --steps 1 --warmup 3 --no-cpu --no-e2e \
    > gpurun_out/ncu_fwd.log 2>&1; echo "ncu fwd rc=$?"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:acoustic_step \
    --launch-skip 60 --launch-count 1 -o gpurun_out/adj_full -f python bench.py --nt 40 --steps 1 --warmup 3 --no-cpu --no-e2e \
    > gpurun_out/ncu_adj.log 2>&1; echo "ncu adj rc=$?"
fi
# small-grid (cluster-resident) kernel: C1 / C5 bench lines and one full capture of a forward + adjoint pair
timeout 300 python bench.py --workload c1 > gpurun_out/bench_c1.log 2>&1; echo "c1 rc=$?"; tail -1 gpurun_out/bench_c1.log | cut -c1-160
timeout 300 python bench.py --workload c5 > gpurun_out/bench_c5.log 2>&1; echo "c5 rc=$?"; tail -1 gpurun_out/bench_c5.log | cut -c1-160
if [ "${NCU:-1}" = "1" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:acoustic_small --launch-skip 6 --launch-count 2 \
    -o gpurun_out/small_full -f python bench.py --workload c1 --nt 200 --steps 1 --warmup 3 --no-cpu --no-e2e \
    > gpurun_out/ncu_small.log 2>&1; echo "ncu small rc=$?"
fi
