#!/bin/bash
# Small-grid (cluster-resident) kernel: the tests that exercise it, then the C1 / f3 / C5 lines
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -k "${PYTEST_K:-acoustic or baseline or edge or dropin or host_kernels or full_size}" > gpurun_out/pytest_small.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_small.log
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_small.log | tail -10
for w in c1 f3 c5; do
  timeout 600 python bench.py --workload $w --no-cpu > gpurun_out/small_$w.log 2>&1
  echo "$w: $(grep -o '"value": [0-9.]*, "unit": "[^"]*"' gpurun_out/small_$w.log | head -1) $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/small_$w.log | head -1)"
done
