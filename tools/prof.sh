#!/bin/bash
# usage (under gpurun): tools_prof.sh <tag> [extra bench args]
#   gpu tests, a plain bench run, then (only after that run exited 0 without ncu) the ncu launch list of the
#   same command and one full capture of forward_culled_kernel.  Outputs in gpurun_out/<tag>_*.
TAG=${1:-r1}; shift
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-stream-leg $*"
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -4 gpurun_out/${TAG}_pytest.log
timeout 300 python bench.py --no-cpu --no-stream-leg $* > gpurun_out/${TAG}_bench.log 2>&1 || { echo "bench failed"; tail -5 gpurun_out/${TAG}_bench.log; exit 1; }
tail -1 gpurun_out/${TAG}_bench.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('value',d['value'],'e2e',d['e2e']['value'],'ms',d['roofline']['ms_per_launch'])"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_l.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:forward_culled -s 4 -c 1 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/${TAG}_ncu_f.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_f.log
