#!/bin/bash
set -u
mkdir -p gpurun_out
for s in 1 2 3 4 5; do
  timeout 300 python bench.py --workload c3 --shots $s --nt 300 --steps 2 --no-cpu --no-e2e > gpurun_out/c3shots$s.log 2>&1
  python - $s gpurun_out/c3shots$s.log <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[2]) if l.startswith("{")][-1]); r=d["roofline"]
print(f"c3 shots {sys.argv[1]}: value {d['value']:7.1f} frac {r['frac']:.4f} fwd {1e3*r['forward']['ms_per_launch']:6.2f} us ({r['forward']['frac']:.3f}) adj {1e3*r['adjoint']['ms_per_launch']:6.2f} us ({r['adjoint']['frac']:.3f})")
PY
done
