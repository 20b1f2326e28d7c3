#!/bin/bash
# 8-GPU lines: shot-parallel default (= the 256-shot survey on 8 GPUs) and the x-slab variant; written to gpurun_out/
set -u
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29521 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_c2_${N}gpu.log 2>&1; echo "c2 rc=$?"; grep '^{' gpurun_out/bench_c2_${N}gpu.log | tail -1 | cut -c1-260
$TR --master-port 29522 bench.py --gpus $N --steps 2 --warmup 3 --workload c4dd > gpurun_out/bench_c4dd_${N}gpu.log 2>&1; echo "c4dd rc=$?"; grep '^{' gpurun_out/bench_c4dd_${N}gpu.log | tail -1 | cut -c1-260
$TR --master-port 29523 bench.py --gpus $N --steps 2 --warmup 3 --workload c5 > gpurun_out/bench_c5_${N}gpu.log 2>&1; echo "c5 rc=$?"; grep '^{' gpurun_out/bench_c5_${N}gpu.log | tail -1 | cut -c1-260
$TR --master-port 29524 bench.py --impl reference --gpus $N --steps 1 --warmup 1 > gpurun_out/bench_reference_${N}gpu.log 2>&1; grep '^{' gpurun_out/bench_reference_${N}gpu.log | tail -1 | cut -c1-200
