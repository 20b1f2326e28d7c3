#!/bin/bash
# Eight-GPU sweep of the x-slab exchange interval (c4dd): gpurun --gpus 8 --timeout 280 -- 'bash scripts/gpu_halo_sweep8.sh'
set -u
mkdir -p gpurun_out
port=29540
for hs in 16 8 12; do
  port=$((port + 1))
  timeout 90 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus 8 --workload c4dd --halo-steps $hs --steps 2 --warmup 3 --no-e2e > gpurun_out/bench_c4dd8_hs$hs.log 2>&1
  echo "halo_steps=$hs rc=$? $(tail -1 gpurun_out/bench_c4dd8_hs$hs.log | cut -c1-160)"
done
