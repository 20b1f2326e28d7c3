#!/bin/bash
# Elastic (C3) tuning: the elastic tests on the default library, then the C3 line per library variant in build/variants/
# usage: bash scripts/gpu_elvariants.sh name1 name2 ...   ("default" = the in-tree library)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "elastic or denise or baseline" > gpurun_out/pytest_el.log 2>&1; echo "pytest rc=$?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_el.log | tail -5
for v in "$@"; do
  if [ "$v" = default ]; then unset FWI_B200_LIB; else export FWI_B200_LIB=$PWD/build/variants/lib_$v.so; fi
  timeout 400 python bench.py --workload c3 --no-cpu --no-e2e > gpurun_out/c3_$v.log 2>&1
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
for l in open(f"gpurun_out/c3_{v}.log"):
    if l.startswith('{"metric'):
        j = json.loads(l); r = j["roofline"]
        print(v, j["value"], "frac", r["frac"], "fwd us", 1e3 * r["forward"]["ms_per_launch"], "adj us", 1e3 * r["adjoint"]["ms_per_launch"])
PY
done
