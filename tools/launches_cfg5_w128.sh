#!/bin/bash
# launch list (ncu, per-kernel durations) of one cfg5 render at w = 128: usage tools/launches_cfg5_w128.sh <points> <tag>
N=${1:-100000000}; TAG=${2:-w128}
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
for _ in range(2):
    out = render_batch(xyz, offs, views, 36, 128, raw=True, need_winner=True, max_item_pts=n)
torch.cuda.synchronize()
PY
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches.csv python /tmp/one_w128.py > gpurun_out/${TAG}_ncu.log 2>&1
python - <<PY
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/${TAG}_launches.csv")) if len(r) > 10 and r[0].isdigit() and "scream::" in r[4]]
half = rows[len(rows)//2:]
agg = collections.OrderedDict()
for r in half:
    k = r[4].split("(")[0]; agg.setdefault(k, [0, 0.0]); agg[k][0] += 1; agg[k][1] += float(r[-1]) / 1e6
tot = sum(v[1] for v in agg.values())
for k, v in agg.items(): print(f"{k:45s} x{v[0]:2d} {v[1]:8.3f} ms")
print("total", round(tot, 3), "ms")
PY
