#!/usr/bin/env python
"""bench.py -- SCREAM scene-renderer hot path on B200 (BASELINE.json metric: depth images/s, Gpoints/s,
% of HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]

Workload (BASELINE.json configs[1]): the renderer's share of one adversarial 3DMatch
training step for B pairs per GPU, n = m = 10,000 points, 6 canonical views, w = 64, rho = 24:
    fake:  forward with winner map (src requires grad)  ->  backward with a N(0,1) upstream gradient
    real:  forward with the GT pose [R|t] fused on the source cloud, no winner map
(train_3d_match.py:186-201).  One step renders B * 2 poses * 6 views * 2 clouds = 24 B depth images.

--impl reference times the reference's CPU implementation (the torch restatement in oracle/, bit-identical
to /root/reference/models/render.py as verified by oracle/make_golden.py; /root/reference does not exist
on the GPU box) on all host cores, each step a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_PTS, V, W_IMG, RHO = 10_000, 6, 64, 24
METRIC, UNIT = "depth_images_per_sec", "images/s"
CLOUD = "unit_ball"


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# --------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# --------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, c in zip(names, r[2:6]) if c.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# --------------------------------------------------------------------------------------------
# synthetic cfg2 batch (host, pinned)
# --------------------------------------------------------------------------------------------
def make_batch(B: int, seed: int, cloud: str = "ball"):
    """B pairs of unit-ball clouds (3DMatch normalisation, three_d_match.py:183-190) + GT poses.
    cloud="shell": the SURVEY 8d realism variant (points on the faces of a random box, sigma 0.003)."""
    from scream_b200 import synth
    rng = np.random.default_rng(seed)
    if cloud == "shell":
        xyz = np.concatenate([synth.room_shell(N_PTS, seed * 7919 + i) for i in range(2 * B)], 0)
    else:
        d = rng.standard_normal((B * 2 * N_PTS, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        xyz = d * (rng.random((B * 2 * N_PTS, 1), dtype=np.float32) ** np.float32(1 / 3))
    pose = np.stack([synth.random_pose(seed * 100003 + b) for b in range(B)])
    offs = np.arange(2 * B + 1, dtype=np.int64) * N_PTS
    grad = rng.standard_normal((B, V, 2, W_IMG, W_IMG), dtype=np.float32)
    return (torch.from_numpy(np.ascontiguousarray(xyz)), torch.from_numpy(offs),
            torch.from_numpy(np.ascontiguousarray(pose)), torch.from_numpy(grad))


def algorithmic_bytes(B: int):
    """SURVEY.md 8d.  fwd: V*12*(n+m) + V*2*w^2*4 (+ the same again for the int32 winner map);
    bwd: V*2*w^2*8 + 12*(n+m)."""
    pts, img = 2 * N_PTS, V * 2 * W_IMG * W_IMG * 4
    fwd_fake = B * (V * 12 * pts + 2 * img)
    fwd_real = B * (V * 12 * pts + img)
    bwd = B * (2 * img + 12 * pts)
    return fwd_fake, fwd_real, bwd


# --------------------------------------------------------------------------------------------
# CPU reference arm / cpu_baseline (the ONLY place bench.py executes oracle/)
# --------------------------------------------------------------------------------------------
def cpu_reference_step(step: int, seed: int = 77):
    """One bounded sample of the workload on the host: ONE pair, ONE of the six views (cycling with the
    step), fake forward+backward and real forward -> 4 depth images.  Cost is exactly linear in views."""
    from oracle import render_oracle as orc
    from scream_b200 import synth
    src = torch.from_numpy(synth.unit_ball(N_PTS, seed))
    tgt = torch.from_numpy(synth.unit_ball(N_PTS, seed + 1))
    pose = torch.from_numpy(synth.random_pose(seed + 2))
    euler = [orc.canonical_eulers("muti")[step % V]]
    g = torch.from_numpy(np.random.default_rng(seed).standard_normal((1, 2, W_IMG, W_IMG)).astype(np.float32))
    t0 = time.perf_counter()
    s = src.clone().requires_grad_(True)
    fake = orc.forward_dense(s, tgt, RHO, W_IMG, euler)            # render.py:61-73, with autograd
    fake.backward(g)
    with torch.no_grad():
        real = orc.forward_dense(orc.apply_pose(pose[:, :3].contiguous(), pose[:, 3:4].contiguous(), src),
                                 tgt, RHO, W_IMG, euler)
    dt = time.perf_counter() - t0
    return dt, 4, float(real.sum())


def run_reference(args, rank: int):
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1: give the reference every host core this process may use
    try:
        avail = len(os.sched_getaffinity(0))
    except AttributeError:
        avail = os.cpu_count() or 1
    if torch.get_num_threads() < avail and "SCREAM_REF_THREADS" not in os.environ:
        torch.set_num_threads(avail)
    elif "SCREAM_REF_THREADS" in os.environ:
        torch.set_num_threads(int(os.environ["SCREAM_REF_THREADS"]))
    threads = torch.get_num_threads()
    budget = 240.0
    t_start = time.perf_counter()
    for i in range(args.warmup):
        cpu_reference_step(i)
        if time.perf_counter() - t_start > budget / 3:
            break
    total, imgs, done = 0.0, 0, 0
    for i in range(args.steps):
        dt, n, _ = cpu_reference_step(i)
        total += dt; imgs += n; done += 1
        if time.perf_counter() - t_start > budget:
            break
    val = imgs / total
    sample = "per step: 1 pair (n=m=10000), 1 of the 6 views, fake fwd+bwd + real fwd = 4 depth images"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": done, "warmup": args.warmup, "ms_per_step": 1e3 * total / done, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(config_dict(args.batch), sampled="1 pair x 1 of the 6 views per step (cost is linear in pairs and views): "
                           "the reference asserts batch 1 (pointnet.py:39-40) and needs N x 4096 temporaries per view"),
            "gpu_launches": 0,
            "gpoints_per_sec": val * N_PTS / 1e9,
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def counter_ball(lo: int, hi: int, seed: int, dev, chunk: int = 8_000_000):
    """Points lo..hi-1 of a unit-ball cloud defined by a COUNTER-based generator (splitmix64 of the global point index, then
    Box-Muller + cube-root radius, all elementwise fp32): point i has the same bits whichever slice it is generated in, so a
    rank's shard IS a slice of the cloud every other world size renders."""
    out = torch.empty((hi - lo, 3), dtype=torch.float32, device=dev)

    def u01(idx, salt):
        z = idx + ((seed * 0x632BE59BD9B4E019 + salt * 0x9E3779B97F4A7C15) & 0x7FFFFFFFFFFFFFFF)
        z = (z ^ (z >> 30)) * -0x40A7B892E31B1A47          # 0xBF58476D1CE4E5B9 as int64
        z = (z ^ (z >> 27)) * -0x6B2FB644ECCEEE15          # 0x94D049BB133111EB as int64
        z = z ^ (z >> 31)
        return (((z >> 20) & 0xFFFFFF).to(torch.float32) + 0.5) * (1.0 / 16777216.0)
    for c0 in range(lo, hi, chunk):
        c1 = min(hi, c0 + chunk)
        idx = torch.arange(c0, c1, dtype=torch.int64, device=dev)
        r1, r2 = torch.sqrt(-2.0 * torch.log(u01(idx, 1))), torch.sqrt(-2.0 * torch.log(u01(idx, 3)))
        a1, a2 = 6.2831853 * u01(idx, 2), 6.2831853 * u01(idx, 4)
        d = torch.stack([r1 * torch.cos(a1), r1 * torch.sin(a1), r2 * torch.cos(a2)], dim=1)
        d = d / d.norm(dim=1, keepdim=True)
        out[c0 - lo:c1 - lo] = d * u01(idx, 5).unsqueeze(1) ** (1.0 / 3.0)
    return out


def image_hash(val: torch.Tensor, winner: torch.Tensor) -> str:
    """64-bit order-sensitive hash of a raw value image and its winner map (bit patterns, wrapping int64 arithmetic)."""
    v = val.contiguous().view(torch.int32).to(torch.int64).reshape(-1)
    w = winner.to(torch.int64).reshape(-1)
    k = torch.arange(1, v.numel() + 1, dtype=torch.int64, device=v.device)
    h = ((v * 1000003 + w) * (2 * k + 1)).sum() + (w * k).sum() * 7919
    return f"{int(h.item()) & 0xFFFFFFFFFFFFFFFF:016x}"


CFG5_SEED_SRC, CFG5_SEED_TGT = 9000, 9500


def streaming_leg_sharded(dev, peak_gbs: float, rank: int, world: int, n: int = 100_000_000, n_views: int = 64, steps: int = 5):
    """The cfg5 cloud of streaming_leg() with its POINTS sharded over the ranks: every rank holds a SLICE of the same
    cloud (counter_ball), the z ranges are combined with all_reduce(MIN/MAX), every rank renders its shard through the
    giant-cloud path and one all_reduce(MAX) over packed (value, index) keys merges the images.  Strong scaling: total work
    fixed.  After the timed region rank 0 renders the WHOLE cloud on its own GPU and compares bit for bit."""
    import torch.distributed as dist
    from scream_b200 import synth
    from scream_b200.render import make_offsets, render_batch, render_cloud_point_sharded
    from scream_b200.shard import balanced_slice
    ns, nt = n // 2, n - n // 2
    (s0, s1), (t0, t1) = balanced_slice(ns, rank, world), balanced_slice(nt, rank, world)
    src, tgt = counter_ball(s0, s1, CFG5_SEED_SRC, dev), counter_ball(t0, t1, CFG5_SEED_TGT, dev)
    views = torch.from_numpy(synth.random_views(n_views, 5)).to(dev)
    xyz_local = torch.cat([src, tgt], 0)          # the shard lives in HBM as one buffer, like the N = 1 cloud
    src, tgt = xyz_local[:s1 - s0], xyz_local[s1 - s0:]
    run = lambda: render_cloud_point_sharded(src, tgt, s0, t0, views, RHO, W_IMG, xyz_local=xyz_local, global_counts=(ns, nt))
    for _ in range(3):
        run()
    dist.barrier(); torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out = run()
    e1.record()
    dist.barrier(); torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    alg = n_views * 12 * n + n_views * 2 * W_IMG * W_IMG * 8
    h = image_hash(out[1], out[2])
    identical = None
    if rank == 0:                                   # the single-GPU render of the same cloud, outside the timed region
        del src, tgt, xyz_local
        torch.cuda.empty_cache()
        xyz = torch.cat([counter_ball(0, ns, CFG5_SEED_SRC, dev), counter_ball(0, nt, CFG5_SEED_TGT, dev)], 0)
        val1, win1, _ = render_batch(xyz, make_offsets([(ns, nt)]).to(dev), views, RHO, W_IMG, raw=True, need_winner=True, max_item_pts=n)
        identical = bool(torch.equal(val1[0], out[1]) and torch.equal(win1[0].to(torch.int64), out[2]))
        h1 = image_hash(val1[0], win1[0])
        del xyz
        torch.cuda.empty_cache()
    return {"workload": f"cfg5: one pair, 100 M points x 64 views, fwd + winner maps, points sharded over {world} GPUs "
                        "(all_reduce of the z ranges, all_reduce(MAX) of the packed key images)",
            "n_points": n, "views": n_views, "ms_per_step": ms, "steps": steps, "scaling": "strong",
            "gpoint_views_per_sec": n_views * n / ms / 1e6, "images_per_sec": 2 * n_views / ms * 1e3,
            "roofline": {"bound": "hbm", "achieved": alg / ms / 1e6, "peak": peak_gbs * world, "unit": "GB/s",
                         "frac": alg / ms / 1e6 / (peak_gbs * world), "algorithmic_bytes_per_step": alg,
                         "note": "ALGORITHMIC throughput, per-view accounting of SURVEY 8d (a kernel that reads a point once for all 64 "
                                 "views moves fewer bytes); peak = measured single-GPU copy bandwidth x GPUs",
                         "frac_of_compulsory_traffic": STREAM_PASSES * 12 * n / ms / 1e6 / (peak_gbs * world)},
            "hash": h, "identical_to_single_gpu": identical, "single_gpu_hash": h1 if rank == 0 else None}


STREAM_PASSES = 8    # times the giant-cloud path reads the cloud from HBM (z-range 4 view groups, seed, box, filter, ...: DESIGN 5.2)


def streaming_leg(dev, peak_gbs: float, n: int = 100_000_000, n_views: int = 64, steps: int = 5):
    """BASELINE.json configs[4] on one GPU (the regime where HBM is the roofline the metric names): one pair of n/2 +
    n/2 unit-ball points (counter_ball: the cloud the N > 1 runs shard), 64 random views, w = 64, rho = 24, forward with
    winner maps, inputs resident in HBM (1.2 GB: ten times the L2).  SURVEY 8d's per-view accounting V*12*N + V*2*w^2*8."""
    from scream_b200 import synth
    from scream_b200.render import render_batch, make_offsets
    ns, nt = n // 2, n - n // 2
    xyz = torch.cat([counter_ball(0, ns, CFG5_SEED_SRC, dev), counter_ball(0, nt, CFG5_SEED_TGT, dev)], 0)
    offs = make_offsets([(ns, nt)]).to(dev)
    views = torch.from_numpy(synth.random_views(n_views, 5)).to(dev)
    run = lambda: render_batch(xyz, offs, views, RHO, W_IMG, raw=True, need_winner=True, max_item_pts=n)
    for _ in range(3):
        run()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out = run()
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / steps
    alg = n_views * 12 * n + n_views * 2 * W_IMG * W_IMG * 8
    res = {"workload": "cfg5: one pair, 100 M points x 64 views, fwd + winner maps, giant-cloud path (exact pre-filter + culled render)",
           "n_points": n, "views": n_views, "ms_per_step": ms, "steps": steps,
           "gpoint_views_per_sec": n_views * n / ms / 1e6, "images_per_sec": 2 * n_views / ms * 1e3,
           "roofline": {"bound": "hbm", "achieved": alg / ms / 1e6, "peak": peak_gbs, "unit": "GB/s",
                        "frac": alg / ms / 1e6 / peak_gbs, "algorithmic_bytes_per_step": alg,
                        "note": "ALGORITHMIC throughput, per-view accounting of SURVEY 8d (a kernel that reads a point once for all 64 "
                                "views moves fewer bytes, so this is not physical HBM utilisation)",
                        "frac_of_compulsory_traffic": STREAM_PASSES * 12 * n / ms / 1e6 / peak_gbs},
           "hash": image_hash(out[0][0], out[1][0])}
    del out
    # cfg5's second shape (rho 36, w 128) on the same cloud: four lattice passes of the pre-filter + the union render
    run2 = lambda: render_batch(xyz, offs, views, 36, 128, raw=True, need_winner=True, max_item_pts=n)
    ms2, _ = _timed(run2, max(2, steps // 2), dev, warm=2)
    alg2 = n_views * 12 * n + n_views * 2 * 128 * 128 * 8
    res["w128_rho36"] = {"ms_per_step": ms2, "gpoint_views_per_sec": n_views * n / ms2 / 1e6, "images_per_sec": 2 * n_views / ms2 * 1e3,
                         "roofline": {"bound": "hbm", "achieved": alg2 / ms2 / 1e6, "peak": peak_gbs, "unit": "GB/s",
                                      "frac": alg2 / ms2 / 1e6 / peak_gbs, "algorithmic_bytes_per_step": alg2,
                                      "note": "per-view accounting, as above"}}
    del xyz
    torch.cuda.empty_cache()
    return res


def _timed(fn, steps, dev, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) / steps, (time.perf_counter() - t0) * 1e3 / steps


def _alg_bytes(B, n, m, V, w, winner=True, bwd=True):
    """forward V*12*(n+m) + V*2*w^2*4 (+ the same again for the winner map); backward V*2*w^2*8 + 12*(n+m)."""
    img = V * 2 * w * w * 4
    return B * (V * 12 * (n + m) + img * (2 if winner else 1)), (B * (2 * img + 12 * (n + m)) if bwd else 0)


def _cpu_sample(src, tgt, rho, w, eulers, backward, block_frac=1.0):
    """One reference render on the host cores (oracle/render_oracle.py, bit-identical to render.py): seconds."""
    from oracle import render_oracle as orc
    s, t = torch.from_numpy(src), torch.from_numpy(tgt)
    t0 = time.perf_counter()
    if backward:
        sg = s.clone().requires_grad_(True)
        out = orc.forward_dense(sg, t, rho, w, eulers)
        out.backward(torch.ones_like(out))
    else:
        with torch.no_grad():
            orc.forward_dense(s, t, rho, w, eulers)
    return time.perf_counter() - t0


def leg_cfg1(dev, peak, cpu: bool):
    """BASELINE configs[0] -- the reference's own call pattern: ONE pair per call through the drop-in module
    (RegistrationRender.forward, pointnet.py:39-40 asserts batch 1), n = m = 10,000 unit-ball points, rho 24, w 64, one
    view ("single depth image") and the six stock views.  Launch-bound, not bandwidth-bound: device and wall time per call."""
    from scream_b200 import synth
    from scream_b200.render import RegistrationRender
    src_h, tgt_h = synth.unit_ball(N_PTS, 4101), synth.unit_ball(N_PTS, 4102)
    src, tgt = torch.from_numpy(src_h).to(dev), torch.from_numpy(tgt_h).to(dev)
    out = {"workload": "cfg1: one 3DMatch-sized pair per call (n = m = 10000), RegistrationRender.forward, w 64, rho 24"}
    for view, V in (("single", 1), ("muti", 6)):
        r = RegistrationRender(RHO, W_IMG, view=view, device=dev)
        g = torch.randn((V, 2, W_IMG, W_IMG), device=dev)

        def fwd():
            with torch.no_grad():
                return r(src, tgt)

        def fwd_bwd():
            s = src.detach().requires_grad_(True)
            r(s, tgt).backward(g)
        f_dev, f_wall = _timed(fwd, 200, dev, warm=20)
        b_dev, b_wall = _timed(fwd_bwd, 200, dev, warm=20)
        # the same with the autograd engine on the calling thread (a user-level switch of torch: no hop to the device's worker thread)
        with torch.autograd.set_multithreading_enabled(False):
            s_dev, s_wall = _timed(fwd_bwd, 200, dev, warm=20)
        fb, bb = _alg_bytes(1, N_PTS, N_PTS, V, W_IMG, winner=False)
        out[f"views_{V}"] = {"fwd_us_device": 1e3 * f_dev, "fwd_us_wall": 1e3 * f_wall, "fwd_bwd_us_device": 1e3 * b_dev,
                             "fwd_bwd_us_wall": 1e3 * b_wall, "fwd_bwd_us_wall_autograd_on_calling_thread": 1e3 * s_wall,
                             "images_per_sec_fwd": 2 * V / (f_dev / 1e3),
                             "roofline": {"bound": "hbm", "achieved": fb / (f_dev / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                                          "frac": fb / (f_dev / 1e3) / 1e9 / peak, "algorithmic_bytes": fb,
                                          "note": "one call = 0.27 MB (V=1): 42 ns at the HBM peak; a call is launch-latency bound"}}
    if cpu:
        from oracle import render_oracle as orc
        _cpu_sample(src_h[:2000], tgt_h[:2000], RHO, W_IMG, orc.canonical_eulers("single"), False)
        dt = _cpu_sample(src_h, tgt_h, RHO, W_IMG, orc.canonical_eulers("single"), False)
        out["cpu_baseline"] = {"value": 2 / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                               "sample": "1 pair, 1 view, forward (the reference's CPU-runnable case)", "ms_per_call": 1e3 * dt}
    return out


def leg_cfg3(dev, peak, cpu: bool, rank: int, world: int, B: int = 64):
    """BASELINE configs[2]: KITTI-scale pairs (n = m = 15,000 bird's-eye slab points), ONE view eulers [0, pi, 0], rho 48
    (train_kitti.py:47-48), batched, B pairs per GPU (items sharded over the ranks, no data-path collective): fake forward
    (+ winner map) + backward + real (posed) forward -> 4 B depth images per step."""
    from scream_b200 import synth
    from scream_b200.render import make_offsets, render_batch, view_matrices
    n = 15_000
    parts = []
    for i in range(B):
        parts += [synth.kitti_slab(n, 5000 + 2 * (rank * B + i)), synth.kitti_slab(n, 5001 + 2 * (rank * B + i))]
    xyz = torch.from_numpy(np.concatenate(parts, 0)).to(dev)
    offs = make_offsets([(n, n)] * B).to(dev)
    pose = torch.from_numpy(np.stack([synth.random_pose(6000 + i) for i in range(B)])).to(dev)
    views = view_matrices([np.array([0, np.pi, 0])]).to(dev)
    grad = torch.randn((B, 1, 2, W_IMG, W_IMG), device=dev)

    def step():
        x = xyz.detach().requires_grad_(True)
        fake, _, _ = render_batch(x, offs, views, 48, W_IMG, need_winner=True, max_item_pts=2 * n)
        fake.backward(grad)
        with torch.no_grad():
            render_batch(xyz, offs, views, 48, W_IMG, pose=pose, need_winner=False, max_item_pts=2 * n)
    ms, _ = _timed(step, 10, dev)
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    f1, b1 = _alg_bytes(B, n, n, 1, W_IMG, winner=True)
    f2, _ = _alg_bytes(B, n, n, 1, W_IMG, winner=False, bwd=False)
    alg = f1 + b1 + f2
    out = {"workload": f"cfg3: KITTI-scale, {B} pairs per GPU x (n = m = 15000 BEV slab), 1 view [0,pi,0], rho 48, w 64: "
                       "fake fwd(+winner) + bwd + real(posed) fwd; items sharded over the GPUs",
           "pairs_per_gpu": B, "n_gpus": world, "ms_per_step": ms, "images_per_sec": 4 * B * world / (ms / 1e3),
           "gpoint_views_per_sec": 2 * 2 * n * B * world / (ms / 1e3) / 1e9,
           "roofline": {"bound": "hbm", "achieved": alg / (ms / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": alg / (ms / 1e3) / 1e9 / peak, "algorithmic_bytes_per_step_per_gpu": alg}}
    if cpu:
        from oracle import render_oracle as orc
        dt = _cpu_sample(parts[0], parts[1], 48, W_IMG, [np.array([0, np.pi, 0])], True)
        out["cpu_baseline"] = {"value": 2 / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                               "sample": "1 pair (15000 + 15000), 1 view, forward + backward", "ms_per_call": 1e3 * dt}
    return out


def leg_cfg4(dev, peak, cpu: bool, B: int = 64):
    """BASELINE configs[3]: OpenGG DSM-DEM patches (DSM = 100 x 100 regular grid: exact ties, dense per-pixel contention;
    DEM = 6 x 6 coarse points), view="single" (pointnet.py:133), at w 64 / 128 / 256 with rho 24 / 36 / 72 (pointnet.py:35:
    rho grows with w); B patches per launch, forward (+ winner map) + backward."""
    from scream_b200 import synth
    from scream_b200.render import make_offsets, render_batch
    parts, sizes = [], []
    for i in range(B):
        dsm, dem = synth.opengg_patch(100, 7000 + i)
        parts += [dsm, dem]; sizes.append((len(dsm), len(dem)))
    xyz = torch.from_numpy(np.concatenate(parts, 0)).to(dev)
    offs = make_offsets(sizes).to(dev)
    views = torch.eye(3, device=dev)[None].contiguous()
    n, m = sizes[0]
    out = {"workload": f"cfg4: {B} OpenGG patches per launch (DSM 100 x 100 grid + DEM 6 x 6), 1 orthographic view, fwd(+winner) + bwd"}
    for w, rho in ((64, 24), (128, 36), (256, 72)):
        grad = torch.randn((B, 1, 2, w, w), device=dev)

        def step():
            x = xyz.detach().requires_grad_(True)
            imgs, _, _ = render_batch(x, offs, views, rho, w, need_winner=True, max_item_pts=n + m)
            imgs.backward(grad)
        ms, _ = _timed(step, 10, dev)
        f, b = _alg_bytes(B, n, m, 1, w, winner=True)
        out[f"w{w}_rho{rho}"] = {"ms_per_step": ms, "images_per_sec": 2 * B / (ms / 1e3), "gpixels_per_sec": 2 * B * w * w / (ms / 1e3) / 1e9,
                                 "roofline": {"bound": "hbm", "achieved": (f + b) / (ms / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                                              "frac": (f + b) / (ms / 1e3) / 1e9 / peak, "algorithmic_bytes_per_step": f + b}}
    if cpu:
        from oracle import render_oracle as orc
        dt = _cpu_sample(parts[0], parts[1], 36, 128, orc.canonical_eulers("single"), False)
        out["cpu_baseline"] = {"value": 2 / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                               "sample": "1 patch, w 128, rho 36, forward", "ms_per_call": 1e3 * dt}
    return out


def leg_disc_stem(dev, peak, B: int = 512):
    """SURVEY 8 f4, the consumer's first layer on the renderer's output (gan.py:39: Conv2d(2, 64, 4, 2, 1) + LeakyReLU(0.2)):
    B x 6 rendered views [B,6,2,64,64] -> [B,6,64,32,32], forward and backward (dx, dW, db).  Bound: the output write (4 * 64
    bytes per output pixel, 8 x the input)."""
    from scream_b200.disc_stem import DiscStem
    torch.manual_seed(0)
    stem = DiscStem(2, 64).to(dev)
    imgs = (torch.rand((B, V, 2, W_IMG, W_IMG), device=dev) * 2 - 1).requires_grad_(True)
    gy = torch.randn((B, V, 64, W_IMG // 2, W_IMG // 2), device=dev)

    def fwd():
        with torch.no_grad():
            return stem(imgs)

    def fwd_bwd():
        imgs.grad = None
        stem.zero_grad(set_to_none=True)
        stem(imgs).backward(gy)
    f_ms, _ = _timed(fwd, 10, dev)
    fb_ms, _ = _timed(fwd_bwd, 10, dev)
    n_out = B * V * 64 * (W_IMG // 2) ** 2 * 4
    n_in = B * V * 2 * W_IMG * W_IMG * 4
    return {"workload": f"f4: discriminator first layer on {B} x 6 rendered views (2 -> 64 channels, k4 s2 p1, LeakyReLU 0.2)",
            "fwd_ms": f_ms, "fwd_bwd_ms": fb_ms, "images_per_sec_fwd": B * V * 2 / (f_ms / 1e3),
            "roofline": {"bound": "hbm", "achieved": (n_in + n_out) / (f_ms / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": (n_in + n_out) / (f_ms / 1e3) / 1e9 / peak, "algorithmic_bytes": n_in + n_out}}


def gpu_eager_baseline(dev):
    """SURVEY 8d's secondary comparator: the reference's own op sequence (oracle/render_oracle.forward_dense = render.py:61-73
    restated op for op) run by stock PyTorch eager ON THE GPU -- what a SCREAM user gets today.  One pair x 6 views, forward."""
    from oracle import render_oracle as orc
    from scream_b200 import synth
    src = torch.from_numpy(synth.unit_ball(N_PTS, 4101)).to(dev)
    tgt = torch.from_numpy(synth.unit_ball(N_PTS, 4102)).to(dev)
    eul = orc.canonical_eulers("muti")

    def fwd():
        with torch.no_grad():
            return orc.forward_dense(src, tgt, RHO, W_IMG, eul)
    try:
        ms, wall = _timed(fwd, 3, dev, warm=1)
    except Exception as e:                                  # noqa: BLE001  (a comparator must not take the bench down)
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}
    torch.cuda.empty_cache()
    return {"workload": "stock PyTorch eager on the B200, the reference's op sequence: 1 pair (10000 + 10000) x 6 views, forward",
            "ms_per_call_device": ms, "ms_per_call_wall": wall, "images_per_sec": 12 / (wall / 1e3)}


def copy_probe(dev, h2d_bytes: int, d2h_bytes: int, chunks: int, streams: int, world: int):
    """The e2e leg's own bound: the same bytes through pinned host memory with NO kernels in between -- one
    cudaMemcpyAsync per chunk buffer, H2D and D2H on `streams` CUDA streams, as the pipeline issues them."""
    h_in = torch.empty(h2d_bytes // chunks, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(d2h_bytes // chunks, dtype=torch.uint8).pin_memory()
    d_in = [torch.empty_like(h_in, device=dev) for _ in range(streams)]
    d_out = [torch.empty(d2h_bytes // chunks, dtype=torch.uint8, device=dev) for _ in range(streams)]
    sts = [torch.cuda.Stream(dev) for _ in range(streams)]

    def step():
        cur = torch.cuda.current_stream(dev)
        for i in range(chunks):
            st = sts[i % streams]
            st.wait_stream(cur)
            with torch.cuda.stream(st):
                d_in[i % streams].copy_(h_in, non_blocking=True)
                h_out.copy_(d_out[i % streams], non_blocking=True)
        for st in sts:
            cur.wait_stream(st)
    ms, _ = _timed(step, 10, dev)
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def config_dict(B: int, world: int = 1, gather: bool = False):
    par = "one GPU" if world == 1 else ("items sharded over the GPUs; every step ends with the all_gather of the depth images and the "
                                        "all_reduce of the statistics (NCCL, overlapped with the next step)" if gather else
                                        "items sharded over the GPUs, no collective in the timed step (--no-gather)")
    return {"workload": "cfg2: 3DMatch training-batch render, fake fwd(+winner)+bwd and real(posed) fwd",
            "pairs_per_gpu": B, "cloud": CLOUD, "n_src": N_PTS, "n_tgt": N_PTS, "views": V, "w": W_IMG, "rho": RHO,
            "images_per_step_per_gpu": 24 * B, "cache": "inputs+outputs per step exceed the 126 MB L2 (B>=512)",
            "parallelism": par}


# --------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------
def bind_to_gpu_numa(index: int):
    """Pin this process (and so its pinned host buffers, first-touch) to the CPU cores NVML reports as local to GPU `index`:
    the e2e leg lives on the host link.  Returns a short description for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = [64 * i + b for i, wd in enumerate(words) for b in range(64) if (wd >> b) & 1]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if allowed:
            os.sched_setaffinity(0, allowed)
            return f"{len(allowed)} cores local to GPU {index} ({allowed[0]}-{allowed[-1]})"
        return "NVML affinity empty: not bound"
    except Exception as e:                                  # noqa: BLE001
        return f"not bound ({type(e).__name__})"


def run_ours(args, rank: int, local_rank: int, world: int):
    import torch.distributed as dist
    affinity = bind_to_gpu_numa(local_rank) if not args.no_bind else "not bound (--no-bind)"
    from scream_b200.render import render_batch, render_pair, view_matrices, RegistrationRender
    from scream_b200 import _lib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    _lib.load()
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    B = args.batch
    views = view_matrices(RegistrationRender(RHO, W_IMG, device="cpu").eulers).to(dev)
    h_xyz, h_offs, h_pose, h_grad = [t.pin_memory() for t in make_batch(B, 1234 + rank, args.cloud)]
    xyz, offs, pose, grad = [t.to(dev) for t in (h_xyz, h_offs, h_pose, h_grad)]
    max_item = 2 * N_PTS
    dense = args.dense
    dual = args.dual and not dense                       # device-resident leg: one dual launch instead of two forwards
    e2e_dual = not dense and not args.no_e2e_dual        # host-buffer leg (StepPipeline): dual launches per chunk by default
    ev = lambda: torch.cuda.Event(enable_timing=True)

    # N > 1: the step ends with the gather north_star names -- every rank receives every rank's depth images (fake and
    # real, [world*B,V,2,w,w] each) and the summed validation statistics.  The collectives are launched asynchronously
    # (NCCL's own stream) and overlap the next step's kernels; all of them are waited for inside the timed region.
    gather = {"on": world > 1 and not args.no_gather, "pending": [], "bufs": None}
    if gather["on"]:
        gather["bufs"] = [[torch.empty((world * B, V, 2, W_IMG, W_IMG), device=dev) for _ in range(2)] for _ in range(2)]

    def gather_step(i, fake, real):
        for h in gather["pending"]:                # at most one step's collectives in flight behind the current one
            h.wait()
        out = gather["bufs"][i % 2]
        stats = torch.stack([fake.detach().sum(dtype=torch.float64), real.sum(dtype=torch.float64)])
        gather["pending"] = [dist.all_gather_into_tensor(out[0], fake.detach(), async_op=True),
                             dist.all_gather_into_tensor(out[1], real, async_op=True),
                             dist.all_reduce(stats, async_op=True)]
        gather["keep"] = (fake, real, stats)

    step_no = [0]

    def step_device(rec=None):
        """fake fwd(+winner) -> bwd -> real fwd, inputs resident in HBM (-> gather of the images at N > 1)."""
        out = step_render(rec)
        if gather["on"]:
            gather_step(step_no[0], out[0], out[1])
            step_no[0] += 1
        return out

    def step_render(rec=None):
        x = xyz.detach().requires_grad_(True)
        if rec is not None:
            e = [ev() for _ in range(6)]; e[0].record()
        if dual:                       # SCREAM_F_DUAL: both renders in one launch, the target cloud sorted once
            fake, real, _, _ = render_pair(x, offs, views, RHO, W_IMG, pose, need_winner=True, max_item_pts=max_item)
            if rec is not None:
                e[1].record(); e[2].record()
            fake.backward(grad)
            if rec is not None:
                e[3].record(); e[4].record(); e[5].record(); rec.append(e)
            return fake, real, x.grad
        fake, _, _ = render_batch(x, offs, views, RHO, W_IMG, dense=dense, need_winner=True, max_item_pts=max_item)
        if rec is not None:
            e[1].record(); e[2].record()
        fake.backward(grad)
        if rec is not None:
            e[3].record(); e[4].record()
        with torch.no_grad():
            real, _, _ = render_batch(xyz, offs, views, RHO, W_IMG, pose=pose, dense=dense, need_winner=False,
                                      max_item_pts=max_item)
        if rec is not None:
            e[5].record(); rec.append(e)
        return fake, real, x.grad

    h_fake = torch.empty((B, V, 2, W_IMG, W_IMG), dtype=torch.float32).pin_memory()
    h_real = torch.empty_like(h_fake).pin_memory()
    h_gxyz = torch.empty_like(h_xyz).pin_memory()

    from scream_b200.pipeline import StepPipeline
    pipe = StepPipeline([(N_PTS, N_PTS)] * B, views, RHO, W_IMG, dev, chunk_pairs=args.chunk, n_streams=args.streams, dense=dense,
                        dual=e2e_dual)

    def step_e2e():
        """Same step through the public host-buffer API (scream_b200.pipeline.StepPipeline): every step copies its
        inputs from pinned host memory and its results back; chunks overlap copies with kernels on CUDA streams."""
        pipe.run(h_xyz, h_pose, h_grad, h_fake, h_real, h_gxyz)
        pipe.join()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, rec=None):
        barrier()
        e0, e1 = ev(), ev()
        e0.record()
        for _ in range(steps):
            fn(rec) if rec is not None else fn()
        for h in gather["pending"]:                # the last step's gather belongs to the timed region
            h.wait()
        gather["pending"] = []
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(args.warmup, 3)):
        step_device()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    rec = []
    ms = timed(step_device, args.steps, rec)
    clocks = sampler.stop() if rank == 0 else None
    ms_nogather, gather_ms = ms, 0.0
    if gather["on"]:                                # the same step without the collectives, and the collectives alone
        gather["on"] = False
        ms_nogather = timed(step_device, args.steps)
        fake, real, _ = step_render()
        bufs = gather["bufs"][0]

        def only_gather():
            dist.all_gather_into_tensor(bufs[0], fake.detach())
            dist.all_gather_into_tensor(bufs[1], real)
        gather_ms = timed(only_gather, args.steps) / args.steps
        gather["on"] = True
    gather["bufs"] = None
    torch.cuda.empty_cache()
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    h2d = int(h_xyz.numel() * 4 + h_pose.numel() * 4 + h_grad.numel() * 4)
    d2h = int(h_fake.numel() * 4 * 2 + h_gxyz.numel() * 4)
    n_chunks = (B + args.chunk - 1) // args.chunk
    copy_ms = copy_probe(dev, h2d, d2h, n_chunks, args.streams, world)

    imgs_step = 24 * B * world
    value = imgs_step * args.steps / (ms / 1e3)
    e2e = imgs_step * args.steps / (ms_e2e / 1e3)
    # per-launch durations of the forward sequence (rank 0, CUDA events on the launching stream)
    t_fake = float(np.mean([e[0].elapsed_time(e[1]) for e in rec]))
    t_bwd = float(np.mean([e[2].elapsed_time(e[3]) for e in rec]))
    t_real = float(np.mean([e[4].elapsed_time(e[5]) for e in rec]))
    b_fake, b_real, b_bwd = algorithmic_bytes(B)
    if dual:        # one launch: the clouds are read once per view, both image sets and both winner maps are written
        b_fake, b_real = b_fake + B * 2 * V * 2 * W_IMG * W_IMG * 4, 0
    peak, peak_src = peaks()
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        t = json.load(open(tpath)).get("forward_culled_kernel_dual" if dual else "forward_culled_kernel", {})
        if t.get("pairs_per_gpu") == B and t.get("cloud") == CLOUD and not dense:
            traffic = t["dram_bytes_read"] + t["dram_bytes_write"]
    ach = (b_fake + b_real) / ((t_fake + t_real) / 1e3) / 1e9
    del pipe
    torch.cuda.empty_cache()
    cfg3 = None if args.no_legs else leg_cfg3(dev, peak, world == 1 and not args.no_cpu and rank == 0, rank, world)
    streaming = None
    if world > 1 and not args.no_stream_leg:           # every rank takes part in the collectives
        streaming = streaming_leg_sharded(dev, peak, rank, world)
    if rank != 0:
        return
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(B, world, gather["on"]),
            "gpoints_per_sec": value * N_PTS / 1e9,
            "gather": None if world == 1 else {
                "in_timed_step": bool(gather["on"]), "what": "all_gather of the fake and the real images to every rank + all_reduce of the statistics, asynchronous (overlaps the next step)",
                "bytes_received_per_rank_per_step": int(2 * world * B * V * 2 * W_IMG * W_IMG * 4),
                "gather_ms": gather_ms, "ms_per_step_without_gather": ms_nogather / args.steps,
                "value_without_gather": imgs_step * args.steps / (ms_nogather / 1e3)},
            "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "copy_roofline_ms": copy_ms, "frac_of_copy_roofline": copy_ms / (ms_e2e / args.steps),
                    "host_affinity": affinity,
                    "copy_roofline": "the same H2D + D2H bytes through pinned memory, one cudaMemcpyAsync per chunk buffer on the same streams, no kernels",
                    "api": f"scream_b200.pipeline.StepPipeline(chunk_pairs={args.chunk}, n_streams={args.streams}{", dual=True" if e2e_dual else ""})"},
            "gpu_launches": args.steps * (4 if dual else 7),
            "kernel": "dense" if dense else ("culled, dual launch (SCREAM_F_DUAL)" if dual else "culled"),
            "roofline": {"bound": "hbm", "kernel": "forward_culled_kernel (timed as one forward call: z-range kernels + it; fake and real calls averaged)",
                         "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": {"fwd_fake": b_fake, "fwd_real": b_real, "bwd": b_bwd},
                         "ms_per_launch": {"fwd_fake": t_fake, "bwd": t_bwd, "fwd_real": t_real}},
            "roofline_bwd": {"bound": "hbm", "kernel": "backward_gather_kernel (one CTA per cloud: the gradient summed in shared memory through the winner map, stored once; grids below 148 CTAs: backward_scatter_kernel, L2 RED.ADD.F32)",
                             "achieved": b_bwd / (t_bwd / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": b_bwd / (t_bwd / 1e3) / 1e9 / peak},
            "clocks": clocks}
    cpu_legs = world == 1 and not args.no_cpu
    if not args.no_legs:
        line["legs"] = {"cfg1": leg_cfg1(dev, peak, cpu_legs), "cfg4": leg_cfg4(dev, peak, cpu_legs),
                        "f4_disc_stem": leg_disc_stem(dev, peak)}
        if world == 1:
            line["gpu_eager_baseline"] = gpu_eager_baseline(dev)
    if cfg3 is not None:
        line.setdefault("legs", {})["cfg3"] = cfg3
    if world == 1 and not args.no_stream_leg:
        line["streaming"] = streaming_leg(dev, peak)
    elif streaming is not None:
        line["streaming"] = streaming
    if world == 1 and not args.no_cpu:
        t0 = time.perf_counter()
        cpu_reference_step(0)                      # warm-up
        tot, n = 0.0, 0
        for i in range(1, 7):
            dt, k, _ = cpu_reference_step(i)
            tot += dt; n += k
            if time.perf_counter() - t0 > 30:
                break
        line["cpu_baseline"] = {"value": n / tot, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                                "sample": f"{n // 4} x (1 pair n=m=10000, 1 view, fake fwd+bwd + real fwd); "
                                          "oracle/render_oracle.py = torch-CPU restatement, bit-identical to reference render.py"}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=512, help="pairs per GPU per step")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dense", action="store_true", help="force the dense exact kernel")
    ap.add_argument("--dual", action="store_true", help="device-resident leg: fake + real render in one launch (SCREAM_F_DUAL, SURVEY 8 f1)")
    ap.add_argument("--no-e2e-dual", action="store_true", help="host-buffer leg: two forward calls per chunk instead of one dual launch")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-bind", action="store_true", help="do not pin the process to the GPU's NUMA-local cores")
    ap.add_argument("--no-legs", action="store_true", help="skip the cfg1 / cfg3 / cfg4 legs and the GPU-eager comparator")
    ap.add_argument("--no-gather", action="store_true", help="N > 1: leave the all_gather of the images out of the timed step")
    ap.add_argument("--no-stream-leg", action="store_true", help="skip the cfg5 streaming-regime leg (one GPU: the whole cloud; N > 1: its points sharded over the ranks)")
    ap.add_argument("--chunk", type=int, default=37, help="pairs per chunk of the host-buffer pipeline (e2e leg)")
    ap.add_argument("--streams", type=int, default=6, help="CUDA streams of the host-buffer pipeline (e2e leg)")
    ap.add_argument("--cloud", default="ball", choices=["ball", "shell"], help="synthetic cloud family (cfg2 default: ball)")
    args = ap.parse_args()
    global CLOUD
    CLOUD = "unit_ball" if args.cloud == "ball" else "room_shell"
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_ours(args, rank, local_rank, world)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.barrier()                              # rank 0 runs the single-GPU legs after the collective ones
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
