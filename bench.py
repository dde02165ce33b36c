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
            "config": config_dict(args.batch), "gpu_launches": 0,
            "gpoints_per_sec": val * N_PTS / 1e9,
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def streaming_leg_sharded(dev, peak_gbs: float, rank: int, world: int, n: int = 100_000_000, n_views: int = 64, steps: int = 5):
    """The same cfg5 cloud with its POINTS sharded over the ranks: every rank holds n / world points, the
    z ranges are combined with all_reduce(MIN/MAX), every rank renders its shard through the giant-cloud path and one
    all_reduce(MAX) over packed (value, index) keys merges the images -- bit-identical to the single-GPU render
    (tests/test_shard_cpu.py, test_point_sharded_merge_equals_unsharded).  Strong scaling: total work fixed."""
    import torch.distributed as dist
    from scream_b200 import synth
    from scream_b200.render import render_cloud_point_sharded
    from scream_b200.shard import balanced_slice
    ns, nt = n // 2, n - n // 2
    (s0, s1), (t0, t1) = balanced_slice(ns, rank, world), balanced_slice(nt, rank, world)

    def ball(count, seed):
        g = torch.Generator(device=dev).manual_seed(seed)
        d = torch.randn((count, 3), generator=g, device=dev)
        d /= d.norm(dim=1, keepdim=True)
        return (d * torch.rand((count, 1), generator=g, device=dev) ** (1.0 / 3.0)).contiguous()
    src, tgt = ball(s1 - s0, 9000 + rank), ball(t1 - t0, 9500 + rank)
    views = torch.from_numpy(synth.random_views(n_views, 5)).to(dev)
    run = lambda: render_cloud_point_sharded(src, tgt, s0, t0, views, RHO, W_IMG)
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
    return {"workload": f"cfg5: one pair, 100 M points x 64 views, fwd + winner maps, points sharded over {world} GPUs "
                        "(all_reduce of the z ranges, all_reduce(MAX) of the packed key images)",
            "n_points": n, "views": n_views, "ms_per_step": ms, "steps": steps, "scaling": "strong",
            "gpoint_views_per_sec": n_views * n / ms / 1e6, "images_per_sec": 2 * n_views / ms * 1e3,
            "roofline": {"bound": "hbm", "achieved": alg / ms / 1e6, "peak": peak_gbs * world, "unit": "GB/s",
                         "frac": alg / ms / 1e6 / (peak_gbs * world), "algorithmic_bytes_per_step": alg,
                         "note": "per-view accounting (SURVEY 8d), peak = measured single-GPU copy bandwidth x GPUs"},
            "checksum": float(out[0].sum())}


def streaming_leg(dev, peak_gbs: float, n: int = 100_000_000, n_views: int = 64, steps: int = 5):
    """BASELINE.json configs[4] on one GPU (the regime where HBM is the roofline the metric names): one pair of n/2 +
    n/2 unit-ball points, 64 random views, w = 64, rho = 24, forward with winner maps, inputs resident in HBM
    (1.2 GB: ten times the L2).  Reported beside the cfg2 headline; SURVEY 8d's per-view accounting V*12*N + V*2*w^2*8."""
    from scream_b200 import synth
    from scream_b200.render import render_batch, make_offsets
    g = torch.Generator(device=dev).manual_seed(1234 + 5)
    d = torch.randn((n, 3), generator=g, device=dev)
    d /= d.norm(dim=1, keepdim=True)
    xyz = (d * torch.rand((n, 1), generator=g, device=dev) ** (1.0 / 3.0)).contiguous()
    del d
    offs = make_offsets([(n // 2, n - n // 2)]).to(dev)
    views = torch.from_numpy(synth.random_views(n_views, 5)).to(dev)
    run = lambda: render_batch(xyz, offs, views, RHO, W_IMG, need_winner=True, max_item_pts=n)
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
                        "note": "per-view accounting (SURVEY 8d); the strict compulsory traffic is 12*N per pass over the cloud"},
           "checksum": float(out[0].sum())}
    del xyz, out
    torch.cuda.empty_cache()
    return res


def config_dict(B: int):
    return {"workload": "cfg2: 3DMatch training-batch render, fake fwd(+winner)+bwd and real(posed) fwd",
            "pairs_per_gpu": B, "cloud": CLOUD, "n_src": N_PTS, "n_tgt": N_PTS, "views": V, "w": W_IMG, "rho": RHO,
            "images_per_step_per_gpu": 24 * B, "cache": "inputs+outputs per step exceed the 126 MB L2 (B>=512)",
            "parallelism": "items sharded over GPUs, no data-path collective"}


# --------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------
def run_ours(args, rank: int, local_rank: int, world: int):
    import torch.distributed as dist
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

    def step_device(rec=None):
        """fake fwd(+winner) -> bwd -> real fwd, inputs resident in HBM."""
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
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)

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
    streaming = None
    if world > 1 and not args.no_stream_leg:           # every rank takes part in the collectives
        del pipe
        torch.cuda.empty_cache()
        streaming = streaming_leg_sharded(dev, peak, rank, world)
    if rank != 0:
        return
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(B),
            "gpoints_per_sec": value * N_PTS / 1e9,
            "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": int(h_xyz.numel() * 4 + h_pose.numel() * 4 + h_grad.numel() * 4),
                    "d2h_bytes_per_step": int(h_fake.numel() * 4 * 2 + h_gxyz.numel() * 4),
                    "api": f"scream_b200.pipeline.StepPipeline(chunk_pairs={args.chunk}, n_streams={args.streams}{", dual=True" if e2e_dual else ""})"},
            "gpu_launches": args.steps * (4 if dual else 7),
            "kernel": "dense" if dense else ("culled, dual launch (SCREAM_F_DUAL)" if dual else "culled"),
            "roofline": {"bound": "hbm", "kernel": "forward_culled_kernel (timed as one forward call: z-range kernels + it; fake and real calls averaged)",
                         "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": {"fwd_fake": b_fake, "fwd_real": b_real, "bwd": b_bwd},
                         "ms_per_launch": {"fwd_fake": t_fake, "bwd": t_bwd, "fwd_real": t_real}},
            "clocks": clocks}
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
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
