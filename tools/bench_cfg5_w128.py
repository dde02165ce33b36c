"""cfg5's second shape (rho 36, w 128) beside w = 64 at the same size: 10^6 (and optionally more) points x 64 views.  From 2^20
points on w = 128 takes the pre-filter as four lattice passes; SCREAM_W128_NOSTREAM=1 times the batch-by-batch culled kernel."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from scream_b200 import synth
from scream_b200.render import render_batch, make_offsets
dev = torch.device("cuda", 0)
views = torch.from_numpy(synth.random_views(64, 5)).to(dev)
for n in [int(a) for a in sys.argv[1:]] or [1_000_000]:
    xyz = torch.cat([bench.counter_ball(0, n // 2, 9000, dev), bench.counter_ball(0, n - n // 2, 9500, dev)], 0)
    offs = make_offsets([(n // 2, n - n // 2)]).to(dev)
    out = {"n_points": n, "views": 64}
    for w, rho in ((64, 24), (128, 36)):
        run = lambda: render_batch(xyz, offs, views, rho, w, raw=True, need_winner=True, max_item_pts=n,
                                   no_stream=(w == 128 and os.environ.get('SCREAM_W128_NOSTREAM') == '1'))
        ms, _ = bench._timed(run, 3, dev, warm=1)
        out[f"w{w}_rho{rho}_ms"] = round(ms, 3)
    print(json.dumps(out), flush=True)
    del xyz
    torch.cuda.empty_cache()
