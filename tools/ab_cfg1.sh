#!/bin/bash
# A/B of compile-time switches on the one-pair call (cfg1) and the cfg2 line.  usage: bash tools/ab_cfg1.sh "-DX=1" ... ""
for defs in "$@"; do
  SCREAM_DEFS="$defs" python -m scream_b200.csrc.build --force --verbose 2>&1 | grep -A2 "21forward_culled" | grep spill
  python tools/bench_cfg1.py 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('[$defs]', {k:round(v['fwd_us_device'],1) for k,v in d.items() if k.startswith('views')})"
  timeout 300 python bench.py --no-cpu --no-stream-leg --no-legs 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('[$defs]', 'cfg2 value',round(d['value']),'ms',{k:round(v,3) for k,v in d['roofline']['ms_per_launch'].items()})"
done
