"""CPU oracle for SCREAM's scene renderer -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module, and only as the checker / the timed CPU baseline.
The product (scream_b200/) never imports it and has no CPU fallback.

This is a restatement (not a copy) of /root/reference/models/render.py in
torch-CPU fp32, written so that every arithmetic step is the same ATen op in the
same order as the reference, which makes it bit-identical to the reference on
the same host (checked by oracle/make_golden.py, which imports the real
reference and asserts equality before it writes tests/golden/).

Parity status: the reference ships no tests or golden vectors for this path
(SURVEY.md section 4), so the oracle is pinned against OUTPUTS OF THE REFERENCE
ITSELF executed in the build container (oracle/make_golden.py -> tests/golden/).

Third-party arithmetic on the path (not under /root/reference):
  torch (pinned 1.7.1 by the reference README, 2.11.0 here): matmul -> sgemm with
    an FMA chain per output element for N >= 15 columns; exp -> ATen vector exp
    (<= 1 ulp, opaque); max(dim) -> first index on ties.
  scipy (pinned 1.5.2, 1.18.1 here): Rotation.from_euler('zyx').as_matrix().
"""
from __future__ import annotations

import math
from typing import Iterable, List, Optional, Sequence, Tuple

import numpy as np
import torch
from scipy.spatial.transform import Rotation

PIXEL_BLOCK = 64 * 64  # render.py:33  (the reference walks the image 4096 pixels at a time)


# --------------------------------------------------------------------------------------
# a1  RegistrationRender.__init__  (render.py:9-25)
# --------------------------------------------------------------------------------------
def pixel_centres(w: int) -> torch.Tensor:
    """[w*w, 2] fp32 pixel centres, (x from column, y from row), render.py:14-16."""
    p = np.arange(w * w)
    col_row = np.stack([p % w, p // w], axis=1)
    grid = torch.from_numpy(col_row).float()
    return (grid - w // 2 + 0.5) / (w // 2)


def canonical_eulers(view: str = "muti") -> List[np.ndarray]:
    """Euler triples (scipy extrinsic 'zyx') of the stock views, render.py:17-25."""
    if view == "muti":
        turn = np.pi
        return [np.array([0, 0, 0]), np.array([0, turn / 2, 0]), np.array([0, turn, 0]),
                np.array([0, turn * 3 / 2, 0]),          # four views about y
                np.array([0, 0, turn / 2]), np.array([0, 0, turn * 3 / 2])]  # two about x
    return [np.array([0, 0, 0])]


# --------------------------------------------------------------------------------------
# a5  scipy euler -> fp32 matrix  (render.py:67)
# --------------------------------------------------------------------------------------
def view_matrices(eulers: Iterable[Sequence[float]]) -> torch.Tensor:
    """[V,3,3] fp32: float64 scipy matrices rounded to fp32 exactly like torch.Tensor(...)."""
    mats = [torch.Tensor(Rotation.from_euler("zyx", e).as_matrix()) for e in eulers]
    return torch.stack(mats, dim=0)


# --------------------------------------------------------------------------------------
# a6  caller-side rigid transform of the source cloud  (train_3d_match.py:198-199)
# --------------------------------------------------------------------------------------
def apply_pose(rot: torch.Tensor, trans: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
    """(R @ src.T + t).T with rot [3,3], trans [3,1] -- sgemm then a separately rounded add."""
    return (torch.matmul(rot, src.t()) + trans).t()


def rotate(view: torch.Tensor, pts: torch.Tensor) -> torch.Tensor:
    """(R @ P.T).T, render.py:68."""
    return torch.matmul(view, pts.T).T


# --------------------------------------------------------------------------------------
# a3  RegistrationRender.render  (render.py:27-59)
# --------------------------------------------------------------------------------------
def depth_weights(x: torch.Tensor) -> Tuple[torch.Tensor, float, float]:
    """pv = 1 - (z - zmin)/(zmax - zmin); zmin/zmax are host scalars, render.py:29-31."""
    z = x[:, 2]
    zmin = torch.min(z, dim=0)[0].item()
    zmax = torch.max(z, dim=0)[0].item()
    pv = 1 - (z - zmin) / (zmax - zmin)
    return pv, zmin, zmax


# --------------------------------------------------------------------------------------
# f2  fp16 (autocast) input mode: what render.py:29-31,68 compute when the caller wraps the generator in
# torch.cuda.amp.autocast() (train_kitti.py:154-169).  Written with EXPLICIT roundings, not with autocast, so it is an
# independent restatement; oracle/make_golden.py checks it bit for bit against the reference executed under
# torch.autocast("cpu", dtype=torch.float16) (the only autocast this container can run).
# --------------------------------------------------------------------------------------
def _h(x: torch.Tensor) -> torch.Tensor:
    """round to fp16, keep as fp32"""
    return x.half().float()


def rotate_half(view: torch.Tensor, pts: torch.Tensor) -> torch.Tensor:
    """render.py:68 under autocast: fp16 operands, fp32 accumulation in k order, fp16 result ([N,3], as fp32)."""
    r, p = _h(view), _h(pts)
    acc = r[:, 0][None] * p[:, 0:1]                       # products of fp16 values are exact in fp32
    acc = acc + r[:, 1][None] * p[:, 1:2]
    acc = acc + r[:, 2][None] * p[:, 2:3]
    return _h(acc)


def depth_weights_half(x: torch.Tensor) -> Tuple[torch.Tensor, float, float]:
    """render.py:29-31 on fp16 depths: every tensor (op) Python-scalar kernel computes in fp32 and rounds to fp16."""
    z = x[:, 2]
    zmin = torch.min(z, dim=0)[0].item()
    zmax = torch.max(z, dim=0)[0].item()
    rng = torch.tensor(zmax - zmin, dtype=torch.float64).float()      # the double difference, cast to the fp32 opmath type
    pv = _h(1 - _h(_h(z - zmin) / rng))
    return pv, zmin, zmax


def _splat_block(xy: torch.Tensor, pv: torch.Tensor, centres: torch.Tensor, rho) -> torch.Tensor:
    """[P, npix] values pv_i * exp(-|xy_i - c_p|^2 / 2 * rho^2), render.py:42-43,46."""
    npix = centres.shape[0]
    d2 = ((xy.view(-1, 1, 2).repeat([1, npix, 1]) - centres.unsqueeze(0)) ** 2).sum(dim=2)
    wgt = torch.exp(-d2 / 2 * rho ** 2)
    return pv.view(-1, 1) * wgt


def render_dense(x: torch.Tensor, src_n: int, rho, w: int,
                 centres: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[2,w,w] un-normalised images of already-rotated points; same op order as render.py:27-59."""
    centres = pixel_centres(w) if centres is None else centres
    pv, _, _ = depth_weights(x)
    rows_src, rows_tgt = [], []
    for blk in range((w // 64) ** 2):
        c = centres[blk * PIXEL_BLOCK:(blk + 1) * PIXEL_BLOCK]
        # the reference evaluates the weights for all N points once, then slices (render.py:42-49)
        wgt = torch.exp(-((x[:, :2].view(-1, 1, 2).repeat([1, PIXEL_BLOCK, 1]) - c.unsqueeze(0)) ** 2)
                        .sum(dim=2) / 2 * rho ** 2)
        rows_src.append(torch.max(pv.view(-1, 1)[:src_n] * wgt[:src_n], dim=0)[0])
        rows_tgt.append(torch.max(pv.view(-1, 1)[src_n:] * wgt[src_n:], dim=0)[0])
    img_s = torch.cat(rows_src, dim=0).view(w, w)
    img_t = torch.cat(rows_tgt, dim=0).view(w, w)
    return torch.stack([img_s, img_t], dim=0)


def forward_dense(src: torch.Tensor, tgt: torch.Tensor, rho, w: int,
                  eulers: Iterable[Sequence[float]]) -> torch.Tensor:
    """[V,2,w,w] in [-1,1): RegistrationRender.forward, render.py:61-73. Differentiable."""
    pts = torch.cat([src, tgt], dim=0)
    centres = pixel_centres(w).to(pts.device)              # render.py:15: pix_xy lives on the module's device
    out = []
    for e in eulers:
        view = torch.Tensor(Rotation.from_euler("zyx", e).as_matrix()).to(pts.device)   # render.py:67
        img = render_dense(rotate(view, pts), src.shape[0], rho, w, centres)
        out.append((img - 0.5) / 0.5)
    return torch.stack(out, dim=0).contiguous()


# --------------------------------------------------------------------------------------
# instrumented, point-chunked variant: values + per-pixel winners + runner-up values.
# Same per-element arithmetic as render_dense; the max over points is merged across
# chunks (first index wins ties, like torch.max(dim=0)).  Bit-identical values.
# --------------------------------------------------------------------------------------
def render_instrumented(x: torch.Tensor, src_n: int, rho, w: int, chunk: int = 2048, half: bool = False,
                        runners: bool = False):
    """Returns dict(val[2,w,w] f32, idx[2,w,w] i64 cloud-local, second[2,w,w] f32, zmin, zmax); runners=True adds the
    near-tie adjudication lists near_pix / near_idx."""
    centres = pixel_centres(w)
    npix = w * w
    pv, zmin, zmax = depth_weights_half(x) if half else depth_weights(x)
    best = torch.full((2, npix), -math.inf)
    second = torch.full((2, npix), -math.inf)
    arg = torch.zeros((2, npix), dtype=torch.int64)
    spans = [(0, src_n), (src_n, x.shape[0])]
    with torch.no_grad():
        for c, (lo, hi) in enumerate(spans):
            for p0 in range(0, npix, PIXEL_BLOCK):
                cen = centres[p0:p0 + PIXEL_BLOCK]
                b = best[c, p0:p0 + PIXEL_BLOCK]
                s = second[c, p0:p0 + PIXEL_BLOCK]
                a = arg[c, p0:p0 + PIXEL_BLOCK]
                for s0 in range(lo, hi, chunk):
                    s1 = min(hi, s0 + chunk)
                    v = _splat_block(x[s0:s1, :2], pv[s0:s1], cen, rho)
                    k = min(2, v.shape[0])
                    top_v, top_i = torch.topk(v, k, dim=0)  # values only; index from max() below
                    cb, ci = torch.max(v, dim=0)             # first index on ties
                    cs = top_v[1] if k == 2 else torch.full_like(cb, -math.inf)
                    take = cb > b                            # strict: earlier chunk keeps ties
                    # runner-up of the union
                    new_second = torch.where(take, torch.maximum(b, cs), torch.maximum(s, cb))
                    s.copy_(new_second)
                    a.copy_(torch.where(take, ci + (s0 - lo), a))
                    b.copy_(torch.where(take, cb, b))
    out = {
        "val": best.view(2, w, w), "idx": arg.view(2, w, w), "second": second.view(2, w, w),
        "zmin": zmin, "zmax": zmax,
    }
    if runners:
        # Near-tie pixels (0 < top-2 gap <= 2 ulp): the points other than the winner whose value reaches the
        # runner-up value, found by re-evaluating the same blocks with the same ATen kernels (so the bits are the bits of
        # the pass above).  [K] flat pixel index into [2, w*w] and [K, 4] cloud-local point indices (-1 padded).
        gap = best - second
        near = (gap > 0) & (gap <= 2 * _ulp(best))
        pix, cand = [], []
        with torch.no_grad():
            for c, (lo, hi) in enumerate(spans):
                for p0 in range(0, npix, PIXEL_BLOCK):
                    cols = near[c, p0:p0 + PIXEL_BLOCK].nonzero().flatten()
                    if cols.numel() == 0:
                        continue
                    found = [[] for _ in range(cols.numel())]
                    for s0 in range(lo, hi, chunk):
                        s1 = min(hi, s0 + chunk)
                        v = _splat_block(x[s0:s1, :2], pv[s0:s1], centres[p0:p0 + PIXEL_BLOCK], rho)[:, cols]
                        hit = (v >= second[c, p0 + cols].unsqueeze(0)).nonzero()
                        for r, k in hit.tolist():
                            found[k].append(r + s0 - lo)
                    for k, col in enumerate(cols.tolist()):
                        others = [i for i in found[k] if i != int(arg[c, p0 + col])][:4]
                        pix.append(c * npix + p0 + col)
                        cand.append(others + [-1] * (4 - len(others)))
        out["near_pix"] = torch.tensor(pix, dtype=torch.int64)
        out["near_idx"] = torch.tensor(cand, dtype=torch.int64).view(-1, 4)
    return out


def _ulp(v: torch.Tensor) -> torch.Tensor:
    """spacing of fp32 values (distance to the next larger magnitude)"""
    a = v.abs()
    return torch.nextafter(a, torch.full_like(a, math.inf)) - a


def forward_instrumented(src: torch.Tensor, tgt: torch.Tensor, rho, w: int, views: torch.Tensor,
                         pose: Optional[torch.Tensor] = None, chunk: int = 2048, half: bool = False,
                         runners: bool = False):
    """Full forward with explicit fp32 view matrices [V,3,3] and optional pose [3,4] on src.
    half=True: the fp16 (autocast) input mode above.

    Returns dict(imgs[V,2,w,w] normalised, val, idx (int32), second, zrange[V,2]).
    """
    if pose is not None:
        src = apply_pose(pose[:, :3].contiguous(), pose[:, 3:4].contiguous(), src)
    pts = torch.cat([src, tgt], dim=0)
    vals, idxs, seconds, zr, npx, nid = [], [], [], [], [], []
    for v in range(views.shape[0]):
        x = rotate_half(views[v], pts) if half else rotate(views[v], pts)
        r = render_instrumented(x, src.shape[0], rho, w, chunk, half=half, runners=runners)
        vals.append(r["val"]); idxs.append(r["idx"]); seconds.append(r["second"])
        zr.append([r["zmin"], r["zmax"]])
        if runners:
            npx.append(r["near_pix"] + v * 2 * w * w); nid.append(r["near_idx"])
    val = torch.stack(vals, dim=0)
    extra = {"near_pix": torch.cat(npx), "near_idx": torch.cat(nid)} if runners else {}
    return {
        **extra,
        "imgs": ((val - 0.5) / 0.5).contiguous(),
        "val": val,
        "idx": torch.stack(idxs, dim=0).to(torch.int32),
        "second": torch.stack(seconds, dim=0),
        "zrange": torch.tensor(zr, dtype=torch.float32),
    }


def pair_values(x: torch.Tensor, pv: torch.Tensor, pts_idx: torch.Tensor, pix_idx: torch.Tensor,
                rho, w: int) -> torch.Tensor:
    """Oracle value of selected (point, pixel) pairs -- used to judge near-tie winners."""
    centres = pixel_centres(w)
    d2 = ((x[pts_idx, :2] - centres[pix_idx]) ** 2).sum(dim=1)
    return pv[pts_idx] * torch.exp(-d2 / 2 * rho ** 2)


# --------------------------------------------------------------------------------------
# a4  backward: autograd through the dense restatement, and the closed form it equals
# --------------------------------------------------------------------------------------
def forward_backward_autograd(src, tgt, rho, w, eulers, grad_out):
    """Reference semantics: autograd through forward_dense.  Returns (imgs, gsrc, gtgt)."""
    s = src.detach().clone().requires_grad_(True)
    t = tgt.detach().clone().requires_grad_(True)
    imgs = forward_dense(s, t, rho, w, eulers)
    imgs.backward(grad_out)
    return imgs.detach(), s.grad.detach(), t.grad.detach()


def backward_closed_form(src, tgt, rho, w, views, winner, zrange, grad_out):
    """Closed-form gradient routed through the winner map, float64 accumulate.

    winner [V,2,w,w] cloud-local indices, zrange [V,2], grad_out [V,2,w,w]. Returns (gsrc, gtgt) f64.
    """
    pts = torch.cat([src, tgt], dim=0).double()
    n = src.shape[0]
    cen = pixel_centres(w).double()
    g_pts = torch.zeros_like(pts)
    for v in range(views.shape[0]):
        R = views[v].double()
        X = pts @ R.T
        zmin, zmax = float(zrange[v, 0]), float(zrange[v, 1])
        for c in range(2):
            i = winner[v, c].reshape(-1).long() + (n if c == 1 else 0)
            g = 2.0 * grad_out[v, c].reshape(-1).double()
            dx = X[i, 0] - cen[:, 0]
            dy = X[i, 1] - cen[:, 1]
            wgt = torch.exp(-(dx * dx + dy * dy) / 2 * rho ** 2)
            pv = 1 - (X[i, 2] - zmin) / (zmax - zmin)
            d_d2 = g * pv * wgt * (-(rho ** 2) / 2)
            gX = torch.stack([2 * dx * d_d2, 2 * dy * d_d2, -(g * wgt) / (zmax - zmin)], dim=1)
            g_pts.index_add_(0, i, gX @ R)
    return g_pts[:n], g_pts[n:]
