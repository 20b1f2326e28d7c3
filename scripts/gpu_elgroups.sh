#!/bin/bash
# C3 line per number of concurrent elastic shot groups (FWI_B200_EL_GROUPS)
set -u
mkdir -p gpurun_out
for g in "$@"; do
  FWI_B200_EL_GROUPS=$g timeout 400 python bench.py --workload c3 --no-cpu --no-e2e > gpurun_out/c3_groups$g.log 2>&1
  python - "$g" <<'PY'
import json, sys
g = sys.argv[1]
for l in open(f"gpurun_out/c3_groups{g}.log"):
    if l.startswith('{"metric'):
        j = json.loads(l); r = j["roofline"]
        print("groups", g, j["value"], "frac", r["frac"], "fwd us", 1e3 * r["forward"]["ms_per_launch"], "adj us", 1e3 * r["adjoint"]["ms_per_launch"])
PY
done
