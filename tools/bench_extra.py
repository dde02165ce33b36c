#!/usr/bin/env python
"""Timings of the widened rows on one B200, CUDA events, JSON lines on stdout:
  fp16 (autocast) mode vs fp32 at a KITTI-shaped batch (train_kitti.py:47-48: one bird's-eye view, rho 48), forward + backward;
  nearest target + Kabsch (pointnet.py:71-74) at n = m = 10 k / 30 k, next to the reference's torch ops on the host cores
  (oracle/registration_oracle.py -- the only use of oracle/ here, as the timed CPU baseline)."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from scream_b200 import synth                                           # noqa: E402
from scream_b200.registration import estimate_transform, nearest_target  # noqa: E402
from scream_b200.render import make_offsets, render_batch              # noqa: E402

DEV = torch.device("cuda", 0)


def timed(fn, steps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def kitti_leg(B=64, n=30000):
    parts = []
    for i in range(B):
        parts += [synth.kitti_slab(n, 2 * i), synth.kitti_slab(n, 2 * i + 1)]
    xyz = torch.from_numpy(np.concatenate(parts, 0)).to(DEV)
    offs = make_offsets([(n, n)] * B).to(DEV)
    from scipy.spatial.transform import Rotation
    views = torch.tensor(Rotation.from_euler("zyx", [0, np.pi, 0]).as_matrix(), dtype=torch.float32)[None].contiguous().to(DEV)
    grad = torch.randn((B, 1, 2, 64, 64), device=DEV)
    out = {"workload": f"cfg3: {B} KITTI-shaped pairs, n = m = {n}, 1 view (BEV), rho 48, w 64, fake forward(+winner) + backward"}
    for half in (False, True):
        def step():
            x = xyz.detach().requires_grad_(True)
            imgs, _, _ = render_batch(x, offs, views, 48, 64, need_winner=True, max_item_pts=2 * n, half=half)
            imgs.backward(grad)
        ms = timed(step)
        out["fp16_mode" if half else "fp32"] = {"ms_per_step": ms, "images_per_sec": 2 * B / (ms / 1e3)}
    return out


def opengg_leg(B=64, side=128, coarse=32):
    """BASELINE config 4: OpenGG DSM-DEM patches (regular grids: dense per-pixel contention, exact ties), one orthographic
    view, at growing image sides; rho grows with w so that a splat stays a few pixels wide (pointnet.py:35: rho 36 at w 128)."""
    parts, sizes = [], []
    for i in range(B):
        dsm, dem = synth.opengg_patch(side, 100 + i, coarse)
        parts += [dsm, dem]; sizes.append((len(dsm), len(dem)))
    xyz = torch.from_numpy(np.concatenate(parts, 0)).to(DEV)
    offs = make_offsets(sizes).to(DEV)
    views = torch.eye(3, device=DEV)[None].contiguous()
    out = {"workload": f"cfg4: {B} OpenGG-shaped patches, DSM {side}x{side} + DEM {coarse}x{coarse} grid points, 1 orthographic view, "
                       "forward(+winner) + backward"}
    for w, rho in ((64, 24), (128, 36), (256, 72), (512, 144)):
        grad = torch.randn((B, 1, 2, w, w), device=DEV)

        def step():
            x = xyz.detach().requires_grad_(True)
            imgs, _, _ = render_batch(x, offs, views, rho, w, need_winner=True, max_item_pts=side * side + coarse * coarse)
            imgs.backward(grad)
        ms = timed(step)
        out[f"w{w}_rho{rho}"] = {"ms_per_step": ms, "images_per_sec": 2 * B / (ms / 1e3), "gpixels_per_sec": 2 * B * w * w / (ms / 1e3) / 1e9}
    return out


def single_pair_leg(n=10000):
    """BASELINE config 1 through the drop-in module: ONE 3DMatch-sized pair per call (the reference asserts batch 1,
    pointnet.py:39-40) -- per-call latency, forward alone and forward + backward, single view and the six views."""
    from scream_b200.render import RegistrationRender
    src = torch.from_numpy(synth.unit_ball(n, 1)).to(DEV)
    tgt = torch.from_numpy(synth.unit_ball(n, 2)).to(DEV)
    out = {"workload": f"cfg1: one pair, n = m = {n}, w 64, rho 24, RegistrationRender.forward per call"}
    for view in ("single", "muti"):
        r = RegistrationRender(24, 64, view=view, device=DEV)
        grad = torch.randn((len(r.eulers), 2, 64, 64), device=DEV)
        with torch.no_grad():
            f = timed(lambda: r(src, tgt), steps=50)

        def step():
            s = src.detach().requires_grad_(True)
            r(s, tgt).backward(grad)
        out[view] = {"views": len(r.eulers), "forward_ms": f, "forward_backward_ms": timed(step, steps=50)}
    return out


def registration_leg(n):
    rng = np.random.default_rng(n)
    src = synth.unit_ball(n, 1)
    pose = synth.random_pose(2)
    moved = (src @ pose[:, :3].T + pose[:, 3]).astype(np.float32)
    tgt = moved[rng.permutation(n)] + rng.normal(0, 0.003, (n, 3)).astype(np.float32)
    pred = moved + rng.normal(0, 0.02, (n, 3)).astype(np.float32)
    a, p, t = (torch.from_numpy(np.ascontiguousarray(v)).to(DEV) for v in (src, pred, tgt))
    ms_nn = timed(lambda: nearest_target(p, t, 1.0))
    ms_all = timed(lambda: estimate_transform(a, p, t, 1.0, 0.075))
    from oracle import registration_oracle as reg
    ca, cp, ct = torch.from_numpy(src), torch.from_numpy(pred), torch.from_numpy(np.ascontiguousarray(tgt))
    reg.estimate_transform(ca, cp, ct, 1.0, 0.075)
    t0 = time.perf_counter()
    k = 0
    while time.perf_counter() - t0 < 5 and k < 10:
        reg.estimate_transform(ca, cp, ct, 1.0, 0.075); k += 1
    cpu_ms = (time.perf_counter() - t0) / k * 1e3
    return {"workload": f"pointnet.py:71-74, n = m = {n}", "nearest_target_ms": ms_nn, "pairs_per_sec": n * n / (ms_nn / 1e3),
            "estimate_transform_ms": ms_all, "cpu_reference_ops_ms": cpu_ms, "cpu_threads": torch.get_num_threads()}


if __name__ == "__main__":
    only = sys.argv[sys.argv.index("--only") + 1] if "--only" in sys.argv else None      # kitti | opengg | single | reg
    if only in (None, "kitti"):
        print(json.dumps(kitti_leg()), flush=True)
    if only in (None, "opengg"):
        print(json.dumps(opengg_leg()), flush=True)
    if only in (None, "single"):
        print(json.dumps(single_pair_leg()), flush=True)
    if only in (None, "reg"):
        for n in (10000, 30000):
            print(json.dumps(registration_leg(n)), flush=True)
