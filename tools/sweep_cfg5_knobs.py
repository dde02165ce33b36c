"""cfg5 at shard sizes: default path vs thin seeds + second round (SCREAM_F_THINSEED), CUDA events."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from scream_b200 import synth
from scream_b200.render import render_batch, make_offsets
dev = torch.device("cuda", 0)
views = torch.from_numpy(synth.random_views(64, 5)).to(dev)
for n in (6_250_000, 12_500_000, 25_000_000, 50_000_000):
    xyz = torch.cat([bench.counter_ball(0, n // 2, 9000, dev), bench.counter_ball(0, n - n // 2, 9500, dev)], 0)
    offs = make_offsets([(n // 2, n - n // 2)]).to(dev)
    res = {}
    for name, kw in (("default", {}), ("thin_seed", {"thin_seed": True})):
        run = lambda: render_batch(xyz, offs, views, 24, 64, raw=True, need_winner=True, max_item_pts=n, **kw)
        ms, _ = bench._timed(run, 5, dev)
        res[name] = round(ms, 3)
    print(n, res, flush=True)
    del xyz
    torch.cuda.empty_cache()
