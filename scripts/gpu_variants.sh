#!/bin/bash
# Times the forward / adjoint step kernels of C2 (6 shots per launch) for tuning variants; one line per variant.
set -u
mkdir -p gpurun_out
run() {   # name, env..., extra bench args after --
  name=$1; shift
  envs=(); while [ $# -gt 0 ] && [ "$1" != "--" ]; do envs+=("$1"); shift; done; [ $# -gt 0 ] && shift
  env "${envs[@]}" timeout 300 python bench.py --one-batch --nt 300 --steps 2 --no-cpu --no-e2e "$@" > gpurun_out/var_$name.log 2>&1
  python - "$name" gpurun_out/var_$name.log <<'PY'
import json, sys
name, path = sys.argv[1], sys.argv[2]
try:
    d = json.loads(open(path).read().strip().splitlines()[-1]); r = d["roofline"]
    print(f"{name:28s} value {d['value']:7.1f}  frac {r['frac']:.4f}  fwd {1e3*r['forward']['ms_per_launch']:6.2f} us ({r['forward']['frac']:.3f})  adj {1e3*r['adjoint']['ms_per_launch']:6.2f} us ({r['adjoint']['frac']:.3f})")
except Exception as e:
    print(f"{name:28s} FAILED {e}: {open(path).read()[-300:]}")
PY
}
run default
run pml0 -- --pml 0
run pieces FWI_B200_PIECES=1
for v in "$@"; do run $v FWI_B200_LIB=$PWD/build/variants/lib_$v.so; done
