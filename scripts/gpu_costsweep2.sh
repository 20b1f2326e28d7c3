#!/bin/bash
# Planner cost sweep under the round-2 defaults (forward per 3-shot group, adjoint as two concurrent groups, paired
# imaging): one 6-shot batch, nt 300.  scripts/gpu_costsweep.sh was the sweep of the 6-shot single-launch form.
set -u
mkdir -p gpurun_out
source <(sed -n '/^run() {/,/^}/p' scripts/gpu_variants.sh)
run default
for b in 2.6 3.0 3.5 4.2; do run "M1_band${b}" FWI_B200_COST_BAND_M1=$b; done
for b in 3.4 4.2 6.0; do run "M0_band${b}" FWI_B200_COST_BAND_M0=$b; done
for b in 2.8 4.0 4.6; do run "M3_band${b}" FWI_B200_COST_BAND_M3=$b; done
for sg in 3.0 9.0; do run "M1_seg${sg}" FWI_B200_COST_SEG_M1=$sg; done
run default2
