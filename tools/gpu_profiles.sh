#!/bin/bash
# round-2 evidence set: launch list of the cfg2 step, --set full captures of the forward kernel and of the atomic paths
# (backward scatter: L2 RED.ADD; z-range: redux + shared/global atomics; stream seed: 64-bit shared atomicMax tiles)
TAG=${1:-r02v}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-stream-leg --no-legs"
C5="python tools/bench_cfg5.py --n 100000000 --steps 1"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo plain failed; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_launches_b512.csv $CMD > /dev/null 2>&1
for k in forward_culled backward_scatter zrange_kernel; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o gpurun_out/prof_${TAG}_$k $CMD > gpurun_out/${TAG}_ncu_$k.log 2>&1; tail -1 gpurun_out/${TAG}_ncu_$k.log
done
$C5 > gpurun_out/${TAG}_plain5.log 2>&1 || { echo plain5 failed; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 30 -c 40 --csv --log-file gpurun_out/${TAG}_launches_cfg5_100M.csv $C5 > /dev/null 2>&1
for k in stream_seed stream_filter; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/prof_${TAG}_$k $C5 > gpurun_out/${TAG}_ncu_$k.log 2>&1; tail -1 gpurun_out/${TAG}_ncu_$k.log
done
