#!/bin/bash
# A/B of builds of the cluster-resident kernel on ONE box.  The scalar propagator reaches the C library through
# _b200fwi_torch.so, which always binds to the in-tree libfwi_b200.so (FWI_B200_LIB only redirects the ctypes
# entry points), so each variant build/variants/lib_<name>.so is copied over the in-tree file before its runs.
# usage: bash scripts/gpu_small_ab.sh head grouped ...   (runs on the GPU box's scratch copy of the tree)
set -u
mkdir -p gpurun_out
for rep in 1 2; do
for v in "$@"; do
  cp build/variants/lib_$v.so physicsbasedfwi2_b200/libfwi_b200.so
  if [ $rep = 1 ]; then
    timeout 600 python -m pytest tests -m gpu -q -x -k "acoustic or baseline or edge" > gpurun_out/pytest_small_$v.log 2>&1; echo "$v pytest rc=$? $(grep -E 'passed|failed' gpurun_out/pytest_small_$v.log | tail -1)"
  fi
  for w in c1 f3 c5; do
    timeout 300 python bench.py --workload $w --no-cpu > gpurun_out/small_${w}_$v.log 2>&1
    echo "$v $w: $(grep -o '"value": [0-9.]*, "unit": "[^"]*"' gpurun_out/small_${w}_$v.log | head -1) $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/small_${w}_$v.log | head -1)"
  done
done
done
