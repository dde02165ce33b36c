#!/bin/bash
# full check: every gpu test, smoke, the default bench line (all legs), the reference arm
TAG=${1:-full}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -3 gpurun_out/${TAG}_pytest.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.log 2>&1; tail -1 gpurun_out/${TAG}_smoke.log
( time timeout 900 python bench.py ) > gpurun_out/${TAG}_bench.log 2> gpurun_out/${TAG}_bench.err; tail -3 gpurun_out/${TAG}_bench.err
tail -1 gpurun_out/${TAG}_bench.log | cut -c1-400
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.log 2>&1; tail -1 gpurun_out/${TAG}_bench_ref.log | cut -c1-300
