#!/bin/bash
# ncu --set full capture of the render of a giant cloud's survivor lists (w = 128: the first lattice launch)
# usage: tools/ncu_giant_render.sh <points> <tag>
N=${1:-100000000}; TAG=${2:-giant}
mkdir -p gpurun_out
cat > /tmp/one_w128.py <<PY
import os, sys, torch
sys.path.insert(0, os.getcwd())
import bench
from scream_b200 import synth
from scream_b200.render import render_batch, make_offsets
dev = torch.device("cuda", 0); n = $N
views = torch.from_numpy(synth.random_views(64, 5)).to(dev)
xyz = torch.cat([bench.counter_ball(0, n // 2, 9000, dev), bench.counter_ball(0, n - n // 2, 9500, dev)], 0)
offs = make_offsets([(n // 2, n - n // 2)]).to(dev)
out = render_batch(xyz, offs, views, 36, 128, raw=True, need_winner=True, max_item_pts=n)
torch.cuda.synchronize()
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:forward_culled -c 1 -f -o gpurun_out/prof_${TAG} python /tmp/one_w128.py > gpurun_out/${TAG}_ncu.log 2>&1; tail -2 gpurun_out/${TAG}_ncu.log
