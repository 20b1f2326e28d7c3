#!/bin/bash
# Small-grid (cluster-resident) kernel: the tests that exercise it, then the C1 / f3 / C5 lines with the halo rows
# through cluster.sync (FWI_B200_SMALL_ASYNC=0) and through st.async + mbarrier (1, the default)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "${PYTEST_K:-acoustic or baseline or edge or dropin or full_size}" > gpurun_out/pytest_small.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_small.log
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_small.log | tail -10
for a in 0 1 0 1; do
  for w in c1 f3 c5; do
    FWI_B200_SMALL_ASYNC=$a timeout 300 python bench.py --workload $w --no-cpu > gpurun_out/small_${w}_async$a.log 2>&1
    echo "async=$a $w: $(grep -o '"value": [0-9.]*, "unit": "[^"]*"' gpurun_out/small_${w}_async$a.log | head -1) $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/small_${w}_async$a.log | head -1)"
  done
done
