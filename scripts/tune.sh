#!/bin/bash
# dev aid: sweep kernel tuning knobs on the C2 workload (run under gpurun)
for cfg in "$@"; do
  env $cfg python bench.py --no-cpu --no-e2e --nt 600 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$cfg', 'shots', d['config']['shots_per_step'], 'value', d['value'], 'fwd_us', r['forward']['ms_per_launch']*1e3, 'adj_us', r['adjoint']['ms_per_launch']*1e3)"
done
