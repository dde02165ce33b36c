#!/usr/bin/env python
"""BASELINE config 5 (scaling sweep): one synthetic pair of N points (n = m = N/2, unit ball), V = 64 random
views, w = 64, rho = 24, forward only (values + winner map), inputs resident in HBM.
    python tools_bench_cfg5.py [--n 1000000] [--views 64] [--steps 5] [--no-stream]
Prints one JSON line: Gpoint-views/s, depth images/s, % of the HBM roofline by SURVEY 8d's per-view accounting
(V*12*N + V*2*w^2*8 bytes)."""
import argparse, json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from scream_b200 import synth
from scream_b200.render import render_batch, make_offsets

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1_000_000); ap.add_argument("--views", type=int, default=64)
ap.add_argument("--steps", type=int, default=5); ap.add_argument("--no-stream", action="store_true")
ap.add_argument("--winner", type=int, default=1)
a = ap.parse_args()
dev = torch.device("cuda:0")
rng = np.random.default_rng(5)
d = rng.standard_normal((a.n, 3)).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True)
xyz = torch.from_numpy(d * (rng.random((a.n, 1), dtype=np.float32) ** np.float32(1 / 3))).to(dev)
offs = make_offsets([(a.n // 2, a.n - a.n // 2)]).to(dev)
views = torch.from_numpy(synth.random_views(a.views, 5)).to(dev)
run = lambda: render_batch(xyz, offs, views, 24, 64, need_winner=bool(a.winner), no_stream=a.no_stream, max_item_pts=a.n)
for _ in range(3): run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.steps): out = run()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.steps
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
alg = a.views * 12 * a.n + a.views * 2 * 64 * 64 * (8 if a.winner else 4)
print(json.dumps({"workload": "cfg5", "n_points": a.n, "views": a.views, "stream": not a.no_stream, "ms": ms,
                  "gpoint_views_per_sec": a.views * a.n / ms / 1e6, "images_per_sec": 2 * a.views / ms * 1e3,
                  "algorithmic_GBps": alg / ms / 1e6, "hbm_frac_per_view_accounting": alg / ms / 1e6 / peak,
                  "compulsory_bytes": 12 * a.n * 3, "checksum": float(out[0].sum())}))
