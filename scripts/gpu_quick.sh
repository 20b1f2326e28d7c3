#!/bin/bash
# quick GPU pass without ncu: parity tests + dev check + short bench
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python scripts/dev_check.py > gpurun_out/dev_check.log 2>&1; tail -12 gpurun_out/dev_check.log
timeout 600 python bench.py --no-cpu ${BENCH_ARGS:-} > gpurun_out/bench.log 2>&1; echo "bench rc=$?" | tee -a gpurun_out/bench.log
tail -3 gpurun_out/bench.log
