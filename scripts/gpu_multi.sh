#!/bin/bash
set -u
mkdir -p gpurun_out
source <(sed -n '/^run() {/,/^}/p' scripts/gpu_variants.sh)
run default
run fwd_multi FWI_B200_MULTI=1
run fwd_multi_2 FWI_B200_MULTI=1 FWI_B200_FWD_SHOTS=2
run fwd2_seq FWI_B200_FWD_SHOTS=2
python bench.py --no-cpu 2>&1 | grep '^{' | tail -1 | cut -c1-220
