#!/bin/bash
# Two-GPU shot-parallel lines of the elastic (c3) and checkpointed (c4) workloads:
#   gpurun --gpus 2 --timeout 600 -- 'bash scripts/gpu_c3c4_2gpu.sh'
set -u
mkdir -p gpurun_out
port=29560
for wl in c3 c4; do
  port=$((port + 1))
  timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus 2 --workload $wl --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_${wl}_2.log 2>&1
  echo "$wl rc=$? $(tail -1 gpurun_out/bench_${wl}_2.log | cut -c1-200)"
done
