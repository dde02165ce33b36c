#!/bin/bash
# one --set full capture of forward_culled_kernel inside the cfg2 bench step (after the same command exited 0 without ncu)
TAG=${1:-n}; shift
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-stream-leg --no-legs $*"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:forward_culled -s 4 -c 1 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/${TAG}_ncu_f.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_f.log
