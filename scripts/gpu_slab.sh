#!/bin/bash
# Two-GPU pass for the x-slab decomposition: gpurun --gpus 2 --timeout 900 -- 'bash scripts/gpu_slab.sh'
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,memory.total --format=csv > gpurun_out/gpu2.txt 2>&1
timeout 600 python -m pytest tests/test_slab_gpu.py -m gpu -x -q > gpurun_out/pytest_slab.log 2>&1; echo "pytest slab rc=$?" | tee -a gpurun_out/pytest_slab.log
tail -5 gpurun_out/pytest_slab.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus 2 --workload c4dd --steps 2 --warmup 3 --no-e2e > gpurun_out/bench_c4dd_2.log 2>&1; echo "c4dd rc=$?"
tail -2 gpurun_out/bench_c4dd_2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_c2_2.log 2>&1; echo "c2x2 rc=$?"
tail -1 gpurun_out/bench_c2_2.log
