#!/usr/bin/env python
"""Where does the giant-cloud pre-filter (forward_stream.cuh) start to pay?  Forward (+winner) of B items of n+m points,
stream path vs SCREAM_F_NOSTREAM, CUDA events.  usage: tools_sweep_stream.py"""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scream_b200 import synth
from scream_b200.render import RegistrationRender, make_offsets, render_batch, view_matrices
DEV = torch.device("cuda", 0)

def timed(fn, steps=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps

six = view_matrices(RegistrationRender(24, 64, device="cpu").eulers).to(DEV)
for kind, rho in (("slab", 48), ("ball", 24)):
    for V in (1, 6):
        views = six[:V].contiguous() if V > 1 else six[2:3].contiguous()
        for n, B in ((30000, 64), (60000, 32), (120000, 16), (250000, 8), (500000, 4), (1000000, 2)):
            gen = synth.kitti_slab if kind == "slab" else synth.unit_ball
            parts = []
            for i in range(B): parts += [gen(n, 2 * i), gen(n, 2 * i + 1)]
            xyz = torch.from_numpy(np.concatenate(parts, 0)).to(DEV)
            offs = make_offsets([(n, n)] * B).to(DEV)
            r = {"cloud": kind, "rho": rho, "V": V, "pts_per_item": 2 * n, "B": B}
            for ns in (False, True):
                r["nostream_ms" if ns else "stream_ms"] = timed(lambda: render_batch(xyz, offs, views, rho, 64, need_winner=True, max_item_pts=2 * n, no_stream=ns))
            print(json.dumps(r), flush=True)
