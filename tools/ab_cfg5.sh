#!/bin/bash
# A/B of compile-time switches on cfg5 (10^8 points x 64 views): rebuild per definition set, exactness test of the giant-cloud
# path, then the timing.  usage: bash tools/ab_cfg5.sh "-DX=1" ... ""   (end with "" so the shipped build stays in place)
for defs in "$@"; do
  SCREAM_DEFS="$defs" python -m scream_b200.csrc.build --force > /dev/null || { echo "build failed: $defs"; continue; }
  timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "giant or sixteen or million or pooled" 2>&1 | tail -1
  for rep in 1 2; do
  timeout 300 python tools/bench_cfg5.py --n 100000000 --steps 5 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('[$defs]', 'cfg5 ms', round(d['ms'],3))"
  done
done
