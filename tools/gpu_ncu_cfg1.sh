#!/bin/bash
TAG=${1:-c1}
mkdir -p gpurun_out
python tools/bench_cfg1.py > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:forward_culled -s 300 -c 1 -f -o gpurun_out/prof_$TAG python tools/bench_cfg1.py > gpurun_out/${TAG}_ncu_f.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_f.log
