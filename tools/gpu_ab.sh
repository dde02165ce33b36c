#!/bin/bash
# A/B on ONE box: rebuild the library with a constant changed and run the short cfg2 bench + the cfg1 latency leg for each
# usage: tools/gpu_ab.sh <file> <sed-pattern-with-@@> <value1> <value2> ...
F=$1; PAT=$2; shift 2
for v in "$@"; do
  sed -i -E "s/${PAT//@@/[0-9]+}/${PAT//@@/$v}/" $F
  python -m scream_b200.csrc.build --force > /dev/null 2>&1 || { echo "build failed for $v"; continue; }
  for rep in 1 2; do
  python bench.py --no-cpu --no-stream-leg --no-legs 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$v', 'value',round(d['value']),'ms',{k: round(x,3) for k,x in d['roofline']['ms_per_launch'].items()})"
  done
  python tools/bench_cfg1.py 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$v', {k: {a: round(b,1) for a,b in d[k].items() if a.endswith('_wall')} for k in ('views_1','views_6')})"
done
