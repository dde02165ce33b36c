#!/bin/bash
# A/B of compile-time switches on cfg5 at shard sizes (one GPU, unsharded clouds of 6.25 / 12.5 / 25 M points).
for defs in "$@"; do
  SCREAM_DEFS="$defs" python -m scream_b200.csrc.build --force > /dev/null || { echo "build failed: $defs"; continue; }
  for n in 6250000 12500000 25000000; do
  timeout 300 python tools/bench_cfg5.py --n $n --steps 5 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('[$defs]', 'n', d['n_points'], 'ms', round(d['ms'],3))"
  done
done
