#!/bin/bash
# The default bench line on N GPUs of one box (the driver's scaling run does the same): gpurun --gpus N -- 'bash scripts/gpu_scaleN.sh N'
set -u
mkdir -p gpurun_out
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_c2_${N}gpu.log 2>&1; echo "c2 rc=$?"; grep '^{' gpurun_out/bench_c2_${N}gpu.log | tail -1 | cut -c1-400
