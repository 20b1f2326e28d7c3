#!/bin/bash
# dev aid: run dev_time.py under several env settings.  usage: sweep.sh "A=1 B=2" "A=2" ...
for cfg in "$@"; do
  echo "== $cfg"
  env $cfg timeout 200 python scripts/dev_time.py ${SHOTS:-6} ${NT:-50} 2 2>&1 | tail -1
done
