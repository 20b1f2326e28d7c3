#!/bin/bash
# Two-GPU sweep of the ghost-exchange interval of the x-slab decomposition (c4dd), after the full GPU test suite:
#   gpurun --gpus 2 --timeout 1200 -- 'bash scripts/gpu_halo_sweep.sh'
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_2gpu.log
tail -3 gpurun_out/pytest_gpu_2gpu.log
port=29520
for hs in 4 8 16 32; do
  port=$((port + 1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus 2 --workload c4dd --halo-steps $hs --steps 2 --warmup 3 --no-e2e > gpurun_out/bench_c4dd_hs$hs.log 2>&1
  echo "halo_steps=$hs rc=$? $(tail -1 gpurun_out/bench_c4dd_hs$hs.log | cut -c1-160)"
done
