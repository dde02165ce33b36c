#!/bin/bash
# A/B of compile-time switches on the GPU box: for each definition set, rebuild the library and run the cfg2 bench line.
# usage: bash tools/ab_build.sh "-DX=1" "-DX=2" ...   (an empty string = the shipped build; the last build stays in place: end with "")
for defs in "$@"; do
  SCREAM_DEFS="$defs" python -m scream_b200.csrc.build --force > /dev/null || { echo "build failed: $defs"; continue; }
  for rep in 1 2; do
  timeout 300 python bench.py --no-cpu --no-stream-leg --no-legs 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('[$defs]', 'value',round(d['value']),'ms',{k:round(v,3) for k,v in d['roofline']['ms_per_launch'].items()})"
  done
done
