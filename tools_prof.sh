#!/bin/bash
# usage: tools_prof.sh <tag> [kernel-regex] [extra bench args]  -- run under gpurun: launch list + one full capture of the top kernel
TAG=${1:-r1}; PAT=${2:-forward_culled}; shift; shift
CMD="python bench.py --steps 2 --warmup 3 --batch 64 --no-cpu $*"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_l_$TAG.log 2>&1
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$PAT -s 4 -c 1 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_f_$TAG.log 2>&1
tail -3 gpurun_out/ncu_f_$TAG.log
