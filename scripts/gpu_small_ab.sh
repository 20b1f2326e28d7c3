#!/bin/bash
# A/B of the cluster-resident kernel: in-tree library against build/variants/lib_<name>.so (args), C1 / f3 / C5 lines
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "acoustic or baseline or edge or dropin or full_size" > gpurun_out/pytest_small.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_small.log | tail -5
for rep in 1 2; do
for v in default "$@"; do
  if [ "$v" = default ]; then unset FWI_B200_LIB; else export FWI_B200_LIB=$PWD/build/variants/lib_$v.so; fi
  for w in c1 f3 c5; do
    timeout 300 python bench.py --workload $w --no-cpu > gpurun_out/small_${w}_$v.log 2>&1
    echo "$v $w: $(grep -o '"value": [0-9.]*, "unit": "[^"]*"' gpurun_out/small_${w}_$v.log | head -1) $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/small_${w}_$v.log | head -1)"
  done
done
done
