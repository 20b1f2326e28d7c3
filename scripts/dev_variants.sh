#!/bin/bash
# Builds tuning variants of libfwi_b200.so (kernel configuration macros of csrc/acoustic.cu) into build/variants/.
# NOTE: FWI_B200_LIB redirects the ctypes entry points only (elastic, misfit, slab, glue).  The scalar propagator goes
# through _b200fwi_torch.so, which binds to the in-tree libfwi_b200.so: to A/B an acoustic variant copy it over the in-tree
# file on the GPU box (scripts/gpu_small_ab.sh does).
# usage: scripts/dev_variants.sh name "-DFWI_CFG_W2=3 -DFWI_CFG_C2=4 -DFWI_CFG_D2=1" [name2 "flags2" ...]
set -u
cd "$(dirname "$0")/.."
mkdir -p build/variants
S=physicsbasedfwi2_b200/csrc
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  ( nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC -shared $flags \
      -o build/variants/lib_$name.so $S/common.cu $S/acoustic.cu $S/acoustic_small.cu $S/elastic.cu $S/misfit.cu $S/host_kernels.cu $S/glue.cu \
      > build/variants/$name.log 2>&1 && echo "built $name" || echo "FAILED $name" ) &
done
wait
