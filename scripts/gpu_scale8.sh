#!/bin/bash
# 8-GPU pass: shot-parallel C2, x-slab C4, end-to-end C5.  gpurun --gpus 8 --timeout 900 -- 'bash scripts/gpu_scale8.sh'
set -u
mkdir -p gpurun_out
N=${N:-8}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $T --master-port 29521 bench.py --gpus $N --steps 3 --warmup 3 --no-e2e > gpurun_out/bench_c2_$N.log 2>&1; echo "c2 rc=$?"; tail -1 gpurun_out/bench_c2_$N.log | cut -c1-160
timeout 300 $T --master-port 29522 bench.py --gpus $N --workload c4dd --steps 2 --warmup 3 --no-e2e > gpurun_out/bench_c4dd_$N.log 2>&1; echo "c4dd rc=$?"; tail -1 gpurun_out/bench_c4dd_$N.log | cut -c1-160
timeout 300 $T --master-port 29523 bench.py --gpus $N --workload c5 --steps 3 --warmup 3 > gpurun_out/bench_c5_$N.log 2>&1; echo "c5 rc=$?"; tail -1 gpurun_out/bench_c5_$N.log | cut -c1-160
