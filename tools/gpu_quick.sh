#!/bin/bash
# quick iteration: parity of the culled kernel against the dense referee / golden vectors, then the cfg2 bench line
TAG=${1:-q}; shift
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_half_gpu.py -m gpu -x -q -k "not sixteen and not million" > gpurun_out/${TAG}_pytest.log 2>&1; tail -3 gpurun_out/${TAG}_pytest.log
timeout 300 python bench.py --no-cpu --no-stream-leg --no-legs $* > gpurun_out/${TAG}_bench.log 2>&1 || { echo "bench failed"; tail -5 gpurun_out/${TAG}_bench.log; exit 1; }
tail -1 gpurun_out/${TAG}_bench.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms',d['roofline']['ms_per_launch'])"
