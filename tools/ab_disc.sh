#!/bin/bash
# A/B of compile-time switches on the discriminator-stem kernels.  usage: bash tools/ab_disc.sh "-DX=1" ... ""
for defs in "$@"; do
  SCREAM_DEFS="$defs" python -m scream_b200.csrc.build --force > /dev/null || { echo "build failed: $defs"; continue; }
  timeout 300 python -m pytest tests/test_disc_stem.py -m gpu -x -q 2>&1 | tail -1
  echo "[$defs] $(python tools/bench_disc_stem.py | cut -c1-110)"
done
