#!/bin/bash
set -u
mkdir -p gpurun_out
source <(sed -n '/^run() {/,/^}/p' scripts/gpu_variants.sh)
# forward (mode 1)
for b in 3.0 3.6 4.2 5.0; do for sd in 0.8 1.0 1.3; do
  run "M1_band${b}_side${sd}" FWI_B200_COST_BAND_M1=$b FWI_B200_COST_SIDE_M1=$sd
done; done
# adjoint: light (mode 0) and pair (mode 3)
for b in 3.0 3.6 4.2 5.0; do for sd in 0.8 1.0 1.3; do
  run "M0_band${b}_side${sd}" FWI_B200_COST_BAND_M0=$b FWI_B200_COST_SIDE_M0=$sd
done; done
for b in 2.6 3.4 4.2 5.0; do for sd in 0.8 1.0 1.3; do
  run "M3_band${b}_side${sd}" FWI_B200_COST_BAND_M3=$b FWI_B200_COST_SIDE_M3=$sd
done; done
