#!/bin/bash
# GPU tests (all failures reported) + kernel timing lines
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --durations=8 -k "${PYTEST_K:-test}" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | tail -30
bash scripts/gpu_variants.sh "$@"
