#!/bin/bash
# A/B of compile-time switches on cfg5 (10^8 points x 64 views), one timing per definition set; the exactness tests of the
# giant-cloud path run for the LAST non-empty set only.  usage: bash tools/ab_cfg5_fast.sh "-DX=1" ... ""
last=""
for defs in "$@"; do [ -n "$defs" ] && last="$defs"; done
for defs in "$@"; do
  SCREAM_DEFS="$defs" python -m scream_b200.csrc.build --force > /dev/null || { echo "build failed: $defs"; continue; }
  timeout 300 python tools/bench_cfg5.py --n 100000000 --steps 10 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('[$defs]', 'cfg5 ms', round(d['ms'],3), 'checksum', d['checksum'])"
  if [ "$defs" == "$last" ]; then
    timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "giant or sixteen or million or pooled" 2>&1 | tail -1
  fi
done
