"""Generate tests/golden/*.npz by EXECUTING THE REFERENCE (build container only).

Run from the repo root:   python oracle/make_golden.py

It imports /root/reference/models/render.py unmodified (the constructor is bypassed
because render.py:12,15 hard-code cuda:0; render()/forward() at render.py:27-73 run
verbatim on CPU tensors), asserts that oracle/render_oracle.py reproduces the
reference BIT FOR BIT (forward values, autograd gradients), and only then freezes the
instrumented outputs (values, winner maps, runner-up values, z ranges, gradients) as
small fixtures.  /root/reference does not exist on the GPU box; the fixtures travel.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from models.render import RegistrationRender as RefRender  # noqa: E402  (the real reference)

from oracle import render_oracle as orc  # noqa: E402
from scream_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def make_reference(rho, w, view="muti"):
    """Reference module on CPU: ctor bypass + restated attribute set-up (render.py:11-25)."""
    r = object.__new__(RefRender)
    nn.Module.__init__(r)
    r.rho, r.w = rho, w
    i, j = np.arange(w * w) // w, np.arange(w * w) % w
    pix = torch.from_numpy(np.concatenate([j.reshape(-1, 1), i.reshape(-1, 1)], axis=1)).float()
    r.pix_xy = (pix - w // 2 + 0.5) / (w // 2)
    r.eulers = orc.canonical_eulers(view)
    return r


def cloud_hash(src, tgt):
    import hashlib
    return hashlib.sha1(np.ascontiguousarray(src).tobytes() + np.ascontiguousarray(tgt).tobytes()).hexdigest()


def freeze(name, src, tgt, rho, w, eulers, pose=None, grad=False, view="muti", full=None, pose44=None):
    src_t, tgt_t = torch.from_numpy(src), torch.from_numpy(tgt)
    ref = make_reference(rho, w, view)
    ref.eulers = [np.asarray(e, dtype=np.float64) for e in eulers]
    # the caller-side rigid transform of the "real" render (train_3d_match.py:198-199)
    if pose44 is not None:
        # the pose as a lie.torch SE3 and ITS transform (lie/torch/se3.py:142-147), executed from the reference
        from lie.torch import SE3
        T = SE3(torch.from_numpy(pose44)[None])                    # liegroupbase.py:18-23: data [1,4,4]
        src_in = T.transform(src_t[None])[0]
        pose = np.ascontiguousarray(pose44[:3, :])
        p = torch.from_numpy(pose)
        assert torch.equal(src_in, (torch.matmul(p[:, :3], src_t.t()) + p[:, 3:4]).t()), \
            f"{name}: SE3.transform differs from the caller-side matmul + trans"
    elif pose is not None:
        p = torch.from_numpy(pose)
        src_in = (torch.matmul(p[:, :3], src_t.t()) + p[:, 3:4]).t()
    else:
        src_in = src_t
    with torch.no_grad():
        want = ref(src_in, tgt_t)                                  # reference, verbatim
        got = orc.forward_dense(src_in, tgt_t, rho, w, ref.eulers)  # restatement
    assert torch.equal(want, got), f"{name}: dense restatement differs from the reference"
    views = orc.view_matrices(ref.eulers)
    ins = orc.forward_instrumented(src_t, tgt_t, rho, w, views,
                                   pose=None if pose is None else torch.from_numpy(pose), runners=True)
    assert torch.equal(want, ins["imgs"]), f"{name}: chunked restatement differs from the reference"
    if full is not None:
        # BASELINE-size fixture: outputs only (the inputs are re-generated from `full`, a seeded recipe, and checked by hash)
        n_max = max(len(src), len(tgt))
        rec = dict(recipe=np.array(full), xyz_sha1=np.array(cloud_hash(src, tgt)), rho=np.float64(rho), w=np.int64(w),
                   eulers=np.stack(ref.eulers).astype(np.float64), views=views.numpy(),
                   val=ins["val"].numpy(), idx=ins["idx"].numpy().astype(np.int16 if n_max < 32768 else np.int32),
                   tie=(ins["second"] == ins["val"]).numpy(), zrange=ins["zrange"].numpy(),
                   near_pix=ins["near_pix"].numpy().astype(np.int32), near_idx=ins["near_idx"].numpy().astype(np.int32))
        if pose is not None:
            rec["pose"] = pose
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
        print(f"wrote {name} (outputs only): V={len(ref.eulers)} n={len(src)} m={len(tgt)} w={w} rho={rho} "
              f"exact-tie px={int(rec['tie'].sum())} near-tie px={len(rec['near_pix'])}")
        return
    # winner indices must reproduce the value they claim
    rec = dict(src=src, tgt=tgt, rho=np.float64(rho), w=np.int64(w),
               eulers=np.stack(ref.eulers).astype(np.float64), views=views.numpy(),
               imgs=want.numpy(), val=ins["val"].numpy(), idx=ins["idx"].numpy(),
               second=ins["second"].numpy(), zrange=ins["zrange"].numpy(),
               near_pix=ins["near_pix"].numpy().astype(np.int32), near_idx=ins["near_idx"].numpy().astype(np.int32))
    if pose is not None:
        rec["pose"] = pose
    if pose44 is not None:
        rec["pose44"] = pose44
    if grad:
        g = torch.from_numpy(np.random.default_rng(7).standard_normal(want.shape).astype(np.float32))
        s = src_in.clone().requires_grad_(True)
        t = tgt_t.clone().requires_grad_(True)
        ref(s, t).backward(g)                                       # reference autograd
        _, gs, gt = orc.forward_backward_autograd(src_in, tgt_t, rho, w, ref.eulers, g)
        assert torch.equal(s.grad, gs) and torch.equal(t.grad, gt), f"{name}: autograd differs"
        cs, ct = orc.backward_closed_form(src_in, tgt_t, rho, w, views, ins["idx"], ins["zrange"], g)
        err = max((cs - gs.double()).abs().max().item(), (ct - gt.double()).abs().max().item())
        scale = max(gs.abs().max().item(), gt.abs().max().item())
        assert err <= 1e-5 * scale, f"{name}: closed-form backward off by {err} (scale {scale})"
        rec.update(grad_out=g.numpy(), grad_src=gs.numpy(), grad_tgt=gt.numpy())
        print(f"  {name}: closed-form vs autograd rel err {err / scale:.2e}")
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    v = rec["val"]
    print(f"wrote {name}: V={len(ref.eulers)} n={len(src)} m={len(tgt)} w={w} rho={rho} "
          f"max={v.max():.4f} zero_px={(v == 0).mean():.3f}")


FULL_RECIPES = {
    # name: (recipe string understood by scream_b200/synth.py:cloud_from_recipe, rho, w, eulers key, posed)
    "full_cfg2_ball": ("ball:10000:2101|ball:10000:2102", 24, 64, "six", False),
    "full_cfg2_shell_real": ("shell:10000:2103|shell:10000:2104", 24, 64, "six", True),
    "full_cfg3_kitti": ("slab:15000:2105|slab:15000:2106", 48, 64, "bev", False),
    "full_cfg4_w64": ("grid:100:2107", 24, 64, "one", False),
    "full_cfg4_w128": ("grid:100:2107", 36, 128, "one", False),
    "full_cfg4_w256": ("grid:100:2107", 72, 256, "one", False),
}


def freeze_full():
    """BASELINE-size fixtures: the reference executed at cfg2 / cfg3 / cfg4 sizes, outputs only."""
    eul = {"six": orc.canonical_eulers("muti"), "one": orc.canonical_eulers("single"), "bev": [np.array([0, np.pi, 0])]}
    for name, (recipe, rho, w, ek, posed) in FULL_RECIPES.items():
        src, tgt = synth.cloud_from_recipe(recipe)
        freeze(name, src, tgt, rho, w, eul[ek], pose=synth.random_pose(2199) if posed else None,
               view="single" if ek == "one" else "muti", full=recipe)


def freeze_extra():
    """a8: pose given as a lie.torch SE3; NaN semantics (render.py:30-31,46: a NaN coordinate poisons every image)."""
    p44 = np.concatenate([synth.random_pose(81), np.array([[0, 0, 0, 1]], np.float32)], 0)
    freeze("cfg2_se3_pose", synth.room_shell(900, 82), synth.unit_ball(800, 83), 24, 64, orc.canonical_eulers("muti"),
           pose44=np.ascontiguousarray(p44))
    src, tgt = synth.unit_ball(300, 84), synth.unit_ball(280, 85)
    tgt[17, 1] = np.float32(np.nan)
    ref = make_reference(24, 64)
    with torch.no_grad():
        want = ref(torch.from_numpy(src), torch.from_numpy(tgt))
    col = torch.full((5, 3), float("nan"))
    assert torch.isnan(want).all() and (torch.max(col, dim=0)[1] == 0).all()     # all-NaN column: NaN at index 0
    np.savez_compressed(os.path.join(OUT, "nan_point.npz"), src=src, tgt=tgt, rho=np.float64(24), w=np.int64(64),
                        eulers=np.stack(ref.eulers).astype(np.float64), views=orc.view_matrices(ref.eulers).numpy(),
                        imgs=want.numpy())
    print("wrote nan_point: every pixel of every view and cloud is NaN in the reference")


def freeze_half(name, src, tgt, rho, w, eulers):
    """the reference under autocast (train_kitti.py:154-169 uses torch.cuda.amp.autocast(); this container
    has no GPU, so the reference runs under torch.autocast("cpu", dtype=torch.float16): same autocast op lists, the
    matmul of render.py:68 in fp16, render.py:31 in fp16, fp32 from render.py:42 on).  Forward: the explicit-rounding
    restatement must equal it bit for bit.  Backward: the reference's own autograd under autocast (it rounds the
    gradient intermediates to fp16 too) is frozen; the product's straight-through fp32 gradient is compared with it
    to a tolerance in the tests."""
    src_t, tgt_t = torch.from_numpy(src), torch.from_numpy(tgt)
    ref = make_reference(rho, w)
    ref.eulers = [np.asarray(e, dtype=np.float64) for e in eulers]
    with torch.no_grad(), torch.autocast("cpu", dtype=torch.float16):
        want = ref(src_t, tgt_t)
    assert want.dtype == torch.float32
    views = orc.view_matrices(ref.eulers)
    ins = orc.forward_instrumented(src_t, tgt_t, rho, w, views, half=True, runners=True)
    assert torch.equal(want, ins["imgs"]), f"{name}: fp16-mode restatement differs from the reference under autocast"
    with torch.no_grad():
        plain = ref(src_t, tgt_t)
    g = torch.from_numpy(np.random.default_rng(7).standard_normal(want.shape).astype(np.float32))
    s = src_t.clone().requires_grad_(True)
    t = tgt_t.clone().requires_grad_(True)
    with torch.autocast("cpu", dtype=torch.float16):
        out = ref(s, t)
    out.backward(g)
    rec = dict(src=src, tgt=tgt, rho=np.float64(rho), w=np.int64(w), eulers=np.stack(ref.eulers).astype(np.float64),
               views=views.numpy(), imgs=want.numpy(), val=ins["val"].numpy(), idx=ins["idx"].numpy(),
               second=ins["second"].numpy(), zrange=ins["zrange"].numpy(),
               near_pix=ins["near_pix"].numpy().astype(np.int32), near_idx=ins["near_idx"].numpy().astype(np.int32),
               grad_out=g.numpy(), grad_src=s.grad.numpy(), grad_tgt=t.grad.numpy())
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(f"wrote {name}: restatement == reference under autocast(fp16); differs from the fp32 render at "
          f"{(plain != want).float().mean().item():.3f} of the pixels (max {(plain - want).abs().max().item():.2e})")


def freeze_half_all():
    freeze_half("half_kitti_bev", synth.kitti_slab(1800, 61), synth.kitti_slab(1700, 62), 48, 64, [np.array([0, np.pi, 0])])
    freeze_half("half_ball_six", synth.unit_ball(900, 63), synth.room_shell(800, 64), 24, 64, orc.canonical_eulers("muti"))


def reference_function(path, name, namespace):
    """Execute ONE unmodified top-level function of a reference file whose module cannot be imported here
    (utils.py imports open3d): its source segment is compiled as is, in a namespace holding its globals."""
    import ast
    src = open(path).read()
    node = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == name)
    exec(compile(ast.get_source_segment(src, node), path, "exec"), namespace)
    return namespace[name]


def freeze_colorize():
    """tests/golden/colorize_strip.npz: the reference's deep_to_img (utils.py:81-91) run on rendered views."""
    from oracle import colorize_oracle as col
    ref_fn = reference_function("/root/reference/utils.py", "deep_to_img", {"torch": torch, "np": np})
    g = np.load(os.path.join(OUT, "cfg2_ball_fake.npz"))
    imgs = g["imgs"].copy()                                   # [6,2,64,64] from the executed reference renderer
    rng = np.random.default_rng(3)
    extra = rng.uniform(-1, 1, (2, 2, 64, 64)).astype(np.float32)
    extra[0, 1, :32] = extra[0, 0, :32]                       # exact ties between the channels
    extra[1, :, :8] = np.float32(-1.0)                        # background in both channels
    imgs = np.concatenate([imgs, extra], 0)
    t = torch.from_numpy(imgs).permute(0, 2, 3, 1)            # train_3d_match.py:212
    want = np.concatenate([ref_fn(t[j]) for j in range(t.shape[0])], axis=1)     # :213, reference code verbatim
    got = col.views_to_strip(imgs)
    assert want.dtype == np.uint8 and np.array_equal(want, got), "colorize restatement differs from the reference"
    np.savez_compressed(os.path.join(OUT, "colorize_strip.npz"), imgs=imgs, strip=want)
    print(f"wrote colorize_strip: imgs {imgs.shape} -> strip {want.shape}, restatement == reference")


def freeze_registration():
    """tests/golden/registration_*.npz: the reference's square_distance / rigid_transform_3d executed verbatim
    on the call pattern of pointnet.py:71-74."""
    from oracle import registration_oracle as reg
    ns = {"torch": torch, "np": np}
    for fn in ("integrate_trans", "square_distance", "rigid_transform_3d"):
        reference_function("/root/reference/utils.py", fn, ns)
    ref_sq, ref_fit = ns["square_distance"], ns["rigid_transform_3d"]
    cases = [("registration_ball", synth.unit_ball(1500, 71), synth.random_pose(72), 0.02, 1.0, 0.075, 1300),
             ("registration_kitti", synth.kitti_slab(2000, 73), synth.random_pose(74), 0.05, 1.7, 0.9, 1800)]
    for name, src, pose, noise, s, thr, m in cases:
        rng = np.random.default_rng(len(name))
        moved = (src @ pose[:, :3].T + pose[:, 3]).astype(np.float32)
        tgt = (moved[rng.permutation(len(src))[:m]] + rng.normal(0, 0.003, (m, 3))).astype(np.float32)
        pred = (moved + rng.normal(0, noise, moved.shape)).astype(np.float32)       # the generator's src_pred
        pred[:7] += np.float32(3.0)                                                   # a few predictions far from every target
        src_t, pred_t, tgt_t = torch.from_numpy(src), torch.from_numpy(pred), torch.from_numpy(tgt)
        dist, ind = ref_sq(pred_t[None] / s, tgt_t[None] / s)[0].min(dim=1)           # pointnet.py:71, reference code
        valid = dist < thr
        T = ref_fit(src_t[None][:, valid], tgt_t[None][:, ind[valid]])[0]            # pointnet.py:72-74, reference code
        T2, d2, i2 = reg.estimate_transform(src_t, pred_t, tgt_t, s, thr)
        assert torch.equal(dist, d2) and torch.equal(ind, i2) and torch.equal(T, T2), f"{name}: restatement differs"
        w = torch.from_numpy(rng.uniform(0, 1, (2, 400)).astype(np.float32))
        A2 = torch.from_numpy(np.stack([src[:400], src[400:800]]))
        B2 = torch.from_numpy(np.stack([moved[:400], -moved[400:800]]))             # item 1: a reflection (det < 0 branch)
        Tw = ref_fit(A2, B2, w.clone(), 0.3)
        assert torch.equal(Tw, reg.rigid_transform_3d(A2, B2, w.clone(), 0.3)), f"{name}: weighted restatement differs"
        np.savez_compressed(os.path.join(OUT, name + ".npz"), src=src, pred=pred, tgt=tgt, s=np.float64(s), thr=np.float64(thr),
                            dist=dist.numpy(), ind=ind.numpy(), T=T.numpy(), pose=pose,
                            wA=A2.numpy(), wB=B2.numpy(), w=w.numpy(), w_thr=np.float64(0.3), Tw=Tw.numpy())
        print(f"wrote {name}: n={len(src)} m={m} valid={int(valid.sum())} det(Tw[1])={np.linalg.det(Tw[1, :3, :3].numpy()):+.3f}")


def freeze_disc():
    """tests/golden/disc_stem.npz: the first two modules of the reference's NLayerDiscriminator (gan.py:39),
    built with input_nc=2, applied to rendered views; forward and the reference's autograd gradients."""
    from models.gan import NLayerDiscriminator, weights_init           # the real reference (needs only torch)
    from oracle import disc_stem_oracle as dso
    torch.manual_seed(5)
    D = NLayerDiscriminator(input_nc=2, n_layers=3, use_actnorm=False, ndf=64).apply(weights_init)   # loss.py:19-24 with 2 channels
    conv, act = D.main[0], D.main[1]
    g = np.load(os.path.join(OUT, "cfg2_ball_fake.npz"))
    x = torch.from_numpy(g["imgs"][[0, 4]].copy()).requires_grad_(True)  # two [2,64,64] views from the executed reference renderer
    y = act(conv(x))
    gy = torch.from_numpy(np.random.default_rng(11).standard_normal(tuple(y.shape)).astype(np.float32))
    y.backward(gy)
    want = dso.disc_stem(x.detach(), conv.weight.detach(), conv.bias.detach())
    assert torch.equal(want, y.detach()), "disc stem restatement differs from the reference modules"
    np.savez_compressed(os.path.join(OUT, "disc_stem.npz"), x=x.detach().numpy(), weight=conv.weight.detach().numpy(),
                        bias=conv.bias.detach().numpy(), y=y.detach().numpy(), grad_out=gy.numpy(), grad_x=x.grad.numpy(),
                        grad_w=conv.weight.grad.numpy(), grad_b=conv.bias.grad.numpy())
    print(f"wrote disc_stem: x {tuple(x.shape)} -> y {tuple(y.shape)}, restatement == reference modules")


def main():
    os.makedirs(OUT, exist_ok=True)
    if "--only-disc" in sys.argv:
        freeze_disc()
        return
    if "--only-registration" in sys.argv:
        freeze_registration()
        return
    if "--only-colorize" in sys.argv:
        freeze_colorize()
        return
    if "--only-half" in sys.argv:
        freeze_half_all()
        return
    if "--only-full" in sys.argv:
        freeze_full()
        return
    if "--only-extra" in sys.argv:
        freeze_extra()
        return
    torch.manual_seed(0)
    six = orc.canonical_eulers("muti")
    one = orc.canonical_eulers("single")
    # cfg1: one 3DMatch-sized pair, single depth image (reduced n for fixture size)
    freeze("cfg1_ball_single", synth.unit_ball(1500, 11), synth.unit_ball(1400, 12), 24, 64, one)
    # cfg2: six canonical views, "fake" (identity pose, with gradients) and "real" (GT pose)
    freeze("cfg2_ball_fake", synth.unit_ball(700, 21), synth.unit_ball(650, 22), 24, 64, six, grad=True)
    freeze("cfg2_shell_real", synth.room_shell(1200, 23), synth.room_shell(1100, 24), 24, 64, six,
           pose=synth.random_pose(25))
    # cfg3: KITTI bird's-eye view [0, pi, 0], rho = 48 (train_kitti.py:47-48)
    freeze("cfg3_kitti_bev", synth.kitti_slab(1800, 31), synth.kitti_slab(1700, 32), 48, 64,
           [np.array([0, np.pi, 0])], grad=True)
    # cfg4: OpenGG regular grid (exact ties), single view, 64 and 128 pixels
    dsm, dem = synth.opengg_patch(48, 41)
    freeze("cfg4_opengg_w64", dsm, dem, 24, 64, one, view="single", grad=True)
    freeze("cfg4_opengg_w128", dsm, dem, 36, 128, one, view="single")
    # sparse cloud: most pixels are far from every point (exact zeros, denormals)
    freeze("sparse_far_pixels", synth.unit_ball(40, 51) * np.float32(0.2), synth.unit_ball(16, 52), 48, 64, six)
    freeze_extra()
    freeze_colorize()
    freeze_half_all()
    freeze_registration()
    freeze_disc()
    freeze_full()


if __name__ == "__main__":
    main()
