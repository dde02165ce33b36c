"""Drop-in scene renderer: the reference's `RegistrationRender` (models/render.py:8-73) on B200.

Same constructor, same mutable attributes (`rho`, `w`, `pix_xy`, `eulers` -- read at CALL time, because
train_kitti.py:47-48 mutates them after construction), same `forward(src_pred, tgt_pcd) -> [V,2,w,w]`
and `render(x, src_n) -> [2,w,w]`, differentiable w.r.t. both clouds, no parameters and no buffers
(`state_dict() == {}`).  Behind it: torch custom ops `scream::render_fwd` / `scream::render_bwd` that
call the C ABI of libscream_render.so (include/scream_render.h) on the current CUDA stream.
There is no CPU path: CPU tensors raise.

`render_batch` is the batched entry point (B pairs, CSR offsets, optional per-item GT pose): one launch
sequence for a whole training batch, which the reference cannot do (pointnet.py:39-40 asserts batch 1).
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
from scipy.spatial.transform import Rotation
from torch import nn

from . import _lib


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _require_cuda(*ts: torch.Tensor) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("scream_b200 renders on CUDA only (no CPU fallback): got a %s tensor" % t.device)


def _stream_ptr(dev: torch.device) -> int:
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    return torch._C._cuda_getCurrentRawStream(idx)          # what torch.cuda.current_stream(dev).cuda_stream returns, 7 us cheaper


class _on_device:
    """`with torch.cuda.device(dev)` only when dev is not already the current device (the context manager costs ~6 us, a
    one-pair call has a budget of ~40)."""
    __slots__ = ("ctx",)

    def __init__(self, dev: torch.device):
        self.ctx = None if dev.index is None or dev.index == torch.cuda.current_device() else torch.cuda.device(dev)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *a):
        if self.ctx is not None:
            self.ctx.__exit__(*a)


_POISON = os.environ.get("SCREAM_POISON", "") not in ("", "0")


def _empty(shape, dtype, dev) -> torch.Tensor:
    """torch.empty for every buffer the path allocates.  SCREAM_POISON=1 (diagnostics): the block is filled with 0xFF bytes
    first (NaN as fp32, -1 as int32), so a kernel that reads a word it never wrote cannot pass the parity tests by luck."""
    t = torch.empty(shape, dtype=dtype, device=dev)
    if _POISON and t.numel():
        t.view(torch.uint8).fill_(0xFF)
    return t


def as_pose(pose, B: int, dev) -> Optional[torch.Tensor]:
    """[B,3,4] fp32 contiguous [R|t] rows on `dev` from what a SCREAM caller has at hand: a [B,3,4] / [B,4,4] (or, for one
    item, [3,4] / [4,4]) tensor, or a lie.torch SE3 -- anything whose `.data` is a [...,4,4] tensor
    (lie/torch/liegroupbase.py:18-23; SE3.transform, lie/torch/se3.py:142-147, reads exactly data[..., :3, :3] and
    data[..., :3, 3:4]).  None stays None."""
    if pose is None:
        return None
    if not isinstance(pose, torch.Tensor):
        data = getattr(pose, "data", None)
        if not isinstance(data, torch.Tensor):
            raise TypeError("pose must be a tensor or an SE3-like object with a .data tensor of shape [...,4,4]")
        pose = data
    if pose.dim() == 2:
        pose = pose.unsqueeze(0)
    if pose.dim() != 3 or pose.shape[0] != B or pose.shape[2] != 4 or pose.shape[1] not in (3, 4):
        raise ValueError(f"pose must be [B,3,4] or [B,4,4] with B={B}, got {tuple(pose.shape)}")
    _require_cuda(pose)
    return pose[:, :3, :].to(device=dev, dtype=torch.float32).contiguous()


# ----------------------------------------------------------------------------------------------
# custom ops
# ----------------------------------------------------------------------------------------------
def _fwd_impl(xyz, offs, pose, views, w, rho, flags, want_winner, max_item_pts, zrange_in=None):
    _require_cuda(xyz, offs, pose, views, zrange_in)
    lib = _lib.load()
    assert xyz.dtype == torch.float32 and xyz.is_contiguous() and xyz.dim() == 2 and xyz.shape[1] == 3
    assert offs.dtype == torch.int64 and offs.is_contiguous() and offs.numel() % 2 == 1
    assert views.dtype == torch.float32 and views.is_contiguous() and views.shape[1:] == (3, 3)
    B, V, total = (offs.numel() - 1) // 2, views.shape[0], xyz.shape[0]
    if pose is not None:
        assert pose.dtype == torch.float32 and pose.is_contiguous() and pose.shape == (B, 3, 4)
    dev = xyz.device
    with torch.cuda.device(dev):
        imgs = _empty((B, V, 2, w, w), torch.float32, dev)
        winner = _empty((B, V, 2, w, w) if want_winner else (0,), torch.int32, dev)
        if zrange_in is not None:                 # point-sharded clouds: the range of the WHOLE cloud (shard.py)
            assert zrange_in.shape == (B, V, 2) and zrange_in.dtype == torch.float32
            zrange = zrange_in.contiguous()       # an input with SCREAM_F_ZRANGE_IN: read, never written
            flags |= _lib.F_ZRANGE_IN
        else:
            zrange = _empty((B, V, 2), torch.float32, dev)
        nws = _lib.workspace_bytes(B, V, w, total)
        ws = _empty((nws,), torch.uint8, dev)
        rc = lib.scream_render_forward(_ptr(xyz), _ptr(offs), total, max_item_pts, _ptr(pose), _ptr(views),
                                       B, V, w, float(rho), flags, _ptr(imgs),
                                       _ptr(winner) if want_winner else None, _ptr(zrange),
                                       _ptr(ws), nws, _stream_ptr(dev))
    _lib.check(rc, "scream_render_forward")
    return imgs, winner, zrange


@torch.library.custom_op("scream::render_fwd", mutates_args=())
def render_fwd(xyz: torch.Tensor, offs: torch.Tensor, pose: Optional[torch.Tensor], views: torch.Tensor,
               w: int, rho: float, flags: int, want_winner: bool, max_item_pts: int,
               zrange_in: Optional[torch.Tensor] = None
               ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    return _fwd_impl(xyz, offs, pose, views, w, rho, flags, want_winner, max_item_pts, zrange_in)


@render_fwd.register_fake
def _(xyz, offs, pose, views, w, rho, flags, want_winner, max_item_pts, zrange_in=None):
    B, V = (offs.numel() - 1) // 2, views.shape[0]
    return (xyz.new_empty((B, V, 2, w, w)),
            xyz.new_empty((B, V, 2, w, w) if want_winner else (0,), dtype=torch.int32),
            xyz.new_empty((B, V, 2)))


def _bwd_impl(xyz, offs, pose, views, w, rho, flags, grad_imgs, winner, zrange, max_item_pts):
    _require_cuda(xyz, offs, pose, views, grad_imgs, winner, zrange)
    lib = _lib.load()
    B, V, total = (offs.numel() - 1) // 2, views.shape[0], xyz.shape[0]
    assert winner.numel() == B * V * 2 * w * w, "forward was run without a winner map"
    grad_imgs = grad_imgs.contiguous().float()
    dev = xyz.device
    with torch.cuda.device(dev):
        grad_xyz = _empty(tuple(xyz.shape), torch.float32, dev)
        rc = lib.scream_render_backward(_ptr(xyz), _ptr(offs), total, max_item_pts, _ptr(pose), _ptr(views),
                                        B, V, w, float(rho), flags, _ptr(grad_imgs), _ptr(winner),
                                        _ptr(zrange), _ptr(grad_xyz), None, 0, _stream_ptr(dev))
    _lib.check(rc, "scream_render_backward")
    return grad_xyz


@torch.library.custom_op("scream::render_bwd", mutates_args=())
def render_bwd(xyz: torch.Tensor, offs: torch.Tensor, pose: Optional[torch.Tensor], views: torch.Tensor,
               w: int, rho: float, flags: int, grad_imgs: torch.Tensor, winner: torch.Tensor,
               zrange: torch.Tensor, max_item_pts: int) -> torch.Tensor:
    return _bwd_impl(xyz, offs, pose, views, w, rho, flags, grad_imgs, winner, zrange, max_item_pts)


@render_bwd.register_fake
def _(xyz, offs, pose, views, w, rho, flags, grad_imgs, winner, zrange, max_item_pts):
    return torch.empty_like(xyz)


def _setup_ctx(ctx, inputs, output):
    xyz, offs, pose, views, w, rho, flags, want_winner, max_item_pts, _zr_in = inputs
    _, winner, zrange = output
    if pose is not None:
        ctx.save_for_backward(xyz, offs, pose, views, winner, zrange)
    else:
        ctx.save_for_backward(xyz, offs, views, winner, zrange)
    ctx.has_pose = pose is not None
    ctx.cfg = (w, rho, flags, max_item_pts)


def _backward(ctx, g_imgs, g_winner, g_zrange):
    if ctx.has_pose:
        xyz, offs, pose, views, winner, zrange = ctx.saved_tensors
    else:
        xyz, offs, views, winner, zrange = ctx.saved_tensors
        pose = None
    w, rho, flags, max_item_pts = ctx.cfg
    g = torch.ops.scream.render_bwd(xyz, offs, pose, views, w, rho, flags, g_imgs, winner, zrange, max_item_pts)
    return g, None, None, None, None, None, None, None, None, None


render_fwd.register_autograd(_backward, setup_context=_setup_ctx)


class _EagerRender(torch.autograd.Function):
    """The same two C-ABI calls as scream::render_fwd / scream::render_bwd without the custom-op dispatcher: in eager mode
    one pair per call (the reference's batch 1, pointnet.py:39-40) is bound by host overhead, not by the kernels.
    render_batch uses this unless torch.compile is tracing (then the registered ops, which have fake kernels)."""

    @staticmethod
    def forward(ctx, xyz, offs, pose, views, w, rho, flags, want_winner, max_item_pts, zrange_in):
        imgs, winner, zrange = _fwd_impl(xyz, offs, pose, views, w, rho, flags, want_winner, max_item_pts, zrange_in)
        ctx.save_for_backward(xyz, offs, views, winner, zrange, *(() if pose is None else (pose,)))
        ctx.cfg = (w, rho, flags, max_item_pts)
        ctx.mark_non_differentiable(winner, zrange)
        return imgs, winner, zrange

    @staticmethod
    def backward(ctx, g_imgs, g_winner, g_zrange):
        xyz, offs, views, winner, zrange, *rest = ctx.saved_tensors
        w, rho, flags, max_item_pts = ctx.cfg
        g = _bwd_impl(xyz, offs, rest[0] if rest else None, views, w, rho, flags, g_imgs, winner, zrange, max_item_pts)
        return g, None, None, None, None, None, None, None, None, None


# ----------------------------------------------------------------------------------------------
# host-side helpers (a1 / a5 of SURVEY.md section 8)
# ----------------------------------------------------------------------------------------------
def view_matrices(eulers: Sequence[Sequence[float]]) -> torch.Tensor:
    """[V,3,3] fp32 on the host, the reference's own conversion (render.py:67)."""
    return torch.stack([torch.Tensor(Rotation.from_euler("zyx", e).as_matrix()) for e in eulers], dim=0)


def make_offsets(sizes: Sequence[Tuple[int, int]]) -> torch.Tensor:
    """CSR offsets [2B+1] int64 (host) for a list of (n_src, n_tgt)."""
    flat = np.asarray(sizes, dtype=np.int64).reshape(-1)
    return torch.from_numpy(np.concatenate([[0], np.cumsum(flat)]).astype(np.int64))


def render_batch(xyz: torch.Tensor, offs: torch.Tensor, views: torch.Tensor, rho, w: int,
                 pose: Optional[torch.Tensor] = None, raw: bool = False, dense: bool = False,
                 need_winner: Optional[bool] = None, max_item_pts: int = 0,
                 zrange_in: Optional[torch.Tensor] = None, no_stream: bool = False, thin_seed: bool = False,
                 half: bool = False):
    """Render B pairs at once.  xyz [total,3] fp32 cuda, offs [2B+1] int64 cuda, views [V,3,3] fp32 cuda,
    pose [B,3,4] / [B,4,4] / a lie.torch SE3 (as_pose) or None; zrange_in [B,V,2]: use this depth range instead of reducing it (point-sharded
    clouds, scream_b200/shard.py).  half=True: the reference's arithmetic under torch.cuda.amp.autocast()
    (train_kitti.py:154-169: fp16 rotation and depth weight); no pose / zrange_in there.
    Returns (imgs [B,V,2,w,w], winner int32 or empty, zrange [B,V,2])."""
    if w < 64 or w % 64:
        raise ValueError("w must be a positive multiple of 64 (render.py:33-34,55)")
    if need_winner is None:
        need_winner = torch.is_grad_enabled() and xyz.requires_grad
    pose = as_pose(pose, (offs.numel() - 1) // 2, xyz.device)
    if half:
        if pose is not None or zrange_in is not None:
            raise ValueError("half=True (autocast mode) takes neither a pose nor a zrange_in (train_kitti.py:193-194 renders "
                             "the posed pair outside autocast)")
        flags = (_lib.F_RAW if raw else 0) | (_lib.F_DENSE if dense else 0)
        return _HalfRender.apply(xyz, offs, views, int(w), float(rho), int(flags), bool(need_winner), int(max_item_pts))
    flags = (_lib.F_RAW if raw else 0) | (_lib.F_DENSE if dense else 0) | (_lib.F_NOSTREAM if no_stream else 0) \
        | (_lib.F_THINSEED if thin_seed else 0)
    if torch.compiler.is_compiling():
        return torch.ops.scream.render_fwd(xyz, offs, pose, views, int(w), float(rho), int(flags),
                                           bool(need_winner), int(max_item_pts), zrange_in)
    return _EagerRender.apply(xyz, offs, pose, views, int(w), float(rho), int(flags), bool(need_winner), int(max_item_pts),
                              zrange_in)


class _DualRender(torch.autograd.Function):
    """One launch for the two renders of a training step (train_3d_match.py:186-201): the pair as given
    ("fake", differentiable) and the pair with the source posed ("real", a constant for autograd, as the caller-side
    transform of the reference is applied to data, not to a prediction).  C ABI: SCREAM_F_DUAL."""

    @staticmethod
    def forward(ctx, xyz, offs, pose, views, w, rho, flags, want_winner, max_item_pts, real_winner):
        _require_cuda(xyz, offs, pose, views)
        lib = _lib.load()
        assert xyz.dtype == torch.float32 and xyz.is_contiguous() and xyz.dim() == 2 and xyz.shape[1] == 3
        assert offs.dtype == torch.int64 and offs.is_contiguous() and offs.numel() % 2 == 1
        assert views.dtype == torch.float32 and views.is_contiguous() and views.shape[1:] == (3, 3)
        B, V, total = (offs.numel() - 1) // 2, views.shape[0], xyz.shape[0]
        assert pose is not None and pose.dtype == torch.float32 and pose.is_contiguous() and pose.shape == (B, 3, 4)
        dev = xyz.device
        with torch.cuda.device(dev):
            imgs = _empty((2, B, V, 2, w, w), torch.float32, dev)
            if want_winner and not real_winner:
                flags |= _lib.F_DUAL_NOWIN1           # winner map of the render as given only: [1,B,V,2,w,w]
            winner = _empty((2 if real_winner else 1, B, V, 2, w, w) if want_winner else (0,), torch.int32, dev)
            zrange = _empty((2, B, V, 2), torch.float32, dev)
            nws = _lib.workspace_bytes(B, V, w, total)
            ws = _empty((nws,), torch.uint8, dev)
            rc = lib.scream_render_forward(_ptr(xyz), _ptr(offs), total, max_item_pts, _ptr(pose), _ptr(views),
                                           B, V, w, float(rho), flags | _lib.F_DUAL, _ptr(imgs),
                                           _ptr(winner) if want_winner else None, _ptr(zrange),
                                           _ptr(ws), nws, _stream_ptr(dev))
        _lib.check(rc, "scream_render_forward (dual)")
        fake, real = imgs[0], imgs[1]
        ctx.save_for_backward(xyz, offs, views, winner, zrange)
        ctx.cfg = (w, rho, flags, max_item_pts, want_winner)
        ctx.mark_non_differentiable(real, winner, zrange)
        return fake, real, winner, zrange

    @staticmethod
    def backward(ctx, g_fake, g_real, g_winner, g_zrange):
        xyz, offs, views, winner, zrange = ctx.saved_tensors
        w, rho, flags, max_item_pts, want_winner = ctx.cfg
        if not want_winner:
            raise RuntimeError("render_pair was run without a winner map (need_winner=False)")
        g = _bwd_impl(xyz, offs, None, views, w, rho, flags & ~_lib.F_DUAL_NOWIN1, g_fake, winner[0], zrange[0], max_item_pts)
        return g, None, None, None, None, None, None, None, None, None


def render_pair(xyz: torch.Tensor, offs: torch.Tensor, views: torch.Tensor, rho, w: int, pose: torch.Tensor,
                raw: bool = False, need_winner: Optional[bool] = None, max_item_pts: int = 0, real_winner: bool = False):
    """The two renders of a training step in ONE launch: (fake [B,V,2,w,w] = the pairs as given, differentiable w.r.t.
    xyz; real [B,V,2,w,w] = source clouds posed by pose [B,3,4], no gradient; winner int32 [1,B,V,2,w,w] (the render as given:
    the one autograd needs; real_winner=True: [2,B,V,2,w,w], the posed render's map as well) or empty;
    zrange [2,B,V,2]).  Bit-identical to render_batch(...) and render_batch(..., pose=pose); the target clouds are
    sorted once.  Not for giant clouds (the C ABI answers SCREAM_E_UNSUPPORTED)."""
    if w < 64 or w % 64:
        raise ValueError("w must be a positive multiple of 64 (render.py:33-34,55)")
    if need_winner is None:
        need_winner = torch.is_grad_enabled() and xyz.requires_grad
    pose = as_pose(pose, (offs.numel() - 1) // 2, xyz.device)
    return _DualRender.apply(xyz, offs, pose, views, int(w), float(rho), int(_lib.F_RAW if raw else 0),
                             bool(need_winner), int(max_item_pts), bool(real_winner))


PAIR_MAX_PTS = 147456          # scream_render_forward_pair's limit (giant pairs take the batched entry point and its pre-filter)


class _PairRender(torch.autograd.Function):
    """RegistrationRender.forward / .render for ONE pair, the reference's call pattern (render.py:61-73, pointnet.py:39-40):
    the two clouds stay two tensors (no torch.cat, no offsets), one kernel launch forward (scream_render_forward_pair), one
    backward (scream_render_backward_pair) that writes the two gradients.  Same bits as render_batch on the pair."""

    @staticmethod
    def run(src, tgt, views, w, rho, flags, want_winner):
        """-> (imgs [V,2,w,w], winner int32 [V,2,w,w] or None, zrange [V,2] or None); no autograd bookkeeping"""
        lib = _lib.load()
        dev = src.device
        V = views.shape[0]
        with _on_device(dev):
            imgs = _empty((V, 2, w, w), torch.float32, dev)
            winner = zrange = None
            if want_winner:                           # one block: the winner map, then the V (zmin, zmax) pairs as raw words
                blk = _empty((V * 2 * w * w + 2 * V,), torch.int32, dev)
                winner, zrange = blk[:V * 2 * w * w].view(V, 2, w, w), blk[V * 2 * w * w:].view(torch.float32).view(V, 2)
            rc = lib.scream_render_forward_pair(src.data_ptr(), src.shape[0], tgt.data_ptr(), tgt.shape[0], None, views.data_ptr(),
                                                V, w, rho, flags, imgs.data_ptr(), _ptr(winner), _ptr(zrange), _stream_ptr(dev))
        if rc:
            _lib.check(rc, "scream_render_forward_pair")
        return imgs, winner, zrange

    @staticmethod
    def forward(ctx, src, tgt, views, w, rho, flags):
        imgs, winner, zrange = _PairRender.run(src, tgt, views, w, rho, flags, True)
        ctx.save_for_backward(src, tgt, views, winner, zrange)
        ctx.cfg = (w, rho, flags)
        return imgs

    @staticmethod
    def backward(ctx, g_imgs):
        src, tgt, views, winner, zrange = ctx.saved_tensors
        w, rho, flags = ctx.cfg
        dev = src.device
        if g_imgs.dtype != torch.float32 or not g_imgs.is_contiguous():
            g_imgs = g_imgs.contiguous().float()
        n, m = src.shape[0], tgt.shape[0]
        with _on_device(dev):
            g = _empty((n + m, 3), torch.float32, dev)             # one block, two gradients
            g_src, g_tgt = g[:n], g[n:]
            rc = _lib.load().scream_render_backward_pair(src.data_ptr(), n, tgt.data_ptr(), m, None,
                                                         views.data_ptr(), views.shape[0], w, rho, flags, g_imgs.data_ptr(),
                                                         winner.data_ptr(), zrange.data_ptr(), g_src.data_ptr(), g_tgt.data_ptr(),
                                                         _stream_ptr(dev))
        if rc:
            _lib.check(rc, "scream_render_backward_pair")
        return g_src, g_tgt, None, None, None, None


class _HalfRender(torch.autograd.Function):
    """fp16 (autocast) input mode, SURVEY 8 f2 (train_kitti.py:154-169): what the reference computes when its generator
    runs under torch.cuda.amp.autocast() -- render.py:68's matmul in fp16 (fp16 operands and result), render.py:31 as a
    chain of fp16 kernels, fp32 from render.py:42 on.  C ABI: scream_half_rotate, then the forward call on B*V
    pre-rotated items with SCREAM_F_HALF; backward = scream_render_backward + scream_half_rotate_backward (the casts
    pass gradients straight through, in fp32: the reference rounds its gradient intermediates to fp16 as well)."""

    @staticmethod
    def forward(ctx, xyz, offs, views, w, rho, flags, want_winner, max_item_pts):
        _require_cuda(xyz, offs, views)
        lib = _lib.load()
        assert xyz.dtype == torch.float32 and xyz.is_contiguous() and xyz.dim() == 2 and xyz.shape[1] == 3
        assert offs.dtype == torch.int64 and offs.is_contiguous() and offs.numel() % 2 == 1
        assert views.dtype == torch.float32 and views.is_contiguous() and views.shape[1:] == (3, 3)
        B, V, total = (offs.numel() - 1) // 2, views.shape[0], xyz.shape[0]
        dev = xyz.device
        with torch.cuda.device(dev):
            xr = _empty((V * total, 3), torch.float32, dev)
            offs_bv = _empty((2 * B * V + 1,), torch.int64, dev)
            rc = lib.scream_half_rotate(_ptr(xyz), _ptr(offs), total, max_item_pts, _ptr(views), B, V,
                                        _ptr(xr), _ptr(offs_bv), _stream_ptr(dev))
            _lib.check(rc, "scream_half_rotate")
            eye = torch.eye(3, dtype=torch.float32, device=dev).unsqueeze(0).contiguous()
        flags |= _lib.F_HALF
        imgs, winner, zrange = _fwd_impl(xr, offs_bv, None, eye, w, rho, flags, want_winner, max_item_pts, None)
        ctx.save_for_backward(xr, offs_bv, eye, offs, views, winner, zrange)
        ctx.cfg = (w, rho, flags, max_item_pts, want_winner, B, V, total)
        winner = winner.view(B, V, 2, w, w) if want_winner else winner
        ctx.mark_non_differentiable(winner, zrange)
        return imgs.view(B, V, 2, w, w), winner, zrange.view(B, V, 2)

    @staticmethod
    def backward(ctx, g_imgs, g_winner, g_zrange):
        xr, offs_bv, eye, offs, views, winner, zrange = ctx.saved_tensors
        w, rho, flags, max_item_pts, want_winner, B, V, total = ctx.cfg
        if not want_winner:
            raise RuntimeError("the fp16-mode render was run without a winner map (need_winner=False)")
        g_xr = _bwd_impl(xr, offs_bv, None, eye, w, rho, flags, g_imgs.contiguous().view(B * V, 1, 2, w, w), winner, zrange,
                         max_item_pts)
        dev = xr.device
        with torch.cuda.device(dev):
            g = _empty((total, 3), torch.float32, dev)
            rc = _lib.load().scream_half_rotate_backward(_ptr(g_xr), _ptr(offs), total, max_item_pts, _ptr(views), B, V,
                                                         _ptr(g), _stream_ptr(dev))
        _lib.check(rc, "scream_half_rotate_backward")
        return g, None, None, None, None, None, None, None


def depth_range(xyz: torch.Tensor, offs: torch.Tensor, views: torch.Tensor, pose: Optional[torch.Tensor] = None,
                max_item_pts: int = 0) -> torch.Tensor:
    """[B,V,2] (zmin, zmax) of the rotated depth over both clouds of every item (render.py:30-31), on the
    device, no host sync.  Used by the point-sharded path (shard.py) before the cross-rank min/max."""
    _require_cuda(xyz, offs, views, pose)
    lib = _lib.load()
    B, V, total = (offs.numel() - 1) // 2, views.shape[0], xyz.shape[0]
    dev = xyz.device
    with torch.cuda.device(dev):
        zr = _empty((B, V, 2), torch.float32, dev)
        nws = (B * V * 2 * 4 + 255) // 256 * 256          # the transform stage only keeps the encoded z range there
        ws = _empty((nws,), torch.uint8, dev)
        rc = lib.scream_render_transform(_ptr(xyz), _ptr(offs), total, int(max_item_pts), _ptr(pose), _ptr(views),
                                         B, V, None, None, _ptr(zr), _ptr(ws), nws, _stream_ptr(dev))
    _lib.check(rc, "scream_render_transform")
    return zr


_OFFS_CACHE = {}


def _pair_offsets(n: int, m: int, dev) -> torch.Tensor:
    key = (n, m, dev)
    hit = _OFFS_CACHE.get(key)
    if hit is None:
        if len(_OFFS_CACHE) > 256:
            _OFFS_CACHE.clear()
        hit = _OFFS_CACHE[key] = torch.tensor([0, n, n + m], dtype=torch.int64).to(dev)
    return hit


def _sharded_stages(xyz, offs, views, rho, w, zr, rank, world, group, reduce_fn=None):
    """scream_render_forward_sharded, stages 1-3 with the two collectives between them (module docstring of shard.py).
    reduce_fn(tensor, op) replaces torch.distributed.all_reduce (tests emulate the ranks on one GPU).
    -> (val [V,2,w,w] un-normalised, winner int32 [V,2,w,w]) of this rank's slice."""
    import ctypes as C
    import torch.distributed as dist
    lib = _lib.load()
    dev = xyz.device
    V, total = views.shape[0], xyz.shape[0]
    if reduce_fn is None:
        reduce_fn = lambda t, op: dist.all_reduce(t, op=op, group=group)
    with _on_device(dev):
        val = _empty((V, 2, w, w), torch.float32, dev)
        win = _empty((V, 2, w, w), torch.int32, dev)
        nws = _lib.workspace_bytes(1, V, w, total)
        ws = _empty((nws,), torch.uint8, dev)
        kp, tp, nk, nt = C.c_void_p(), C.c_void_p(), C.c_size_t(), C.c_size_t()
        zr1 = zr[None].contiguous()

        def stage(i):
            rc = lib.scream_render_forward_sharded(xyz.data_ptr(), offs.data_ptr(), total, total, views.data_ptr(), 1, V, w,
                                                   float(rho), _lib.F_RAW, i, rank, world, val.data_ptr(), win.data_ptr(),
                                                   zr1.data_ptr(), ws.data_ptr(), nws, C.byref(kp), C.byref(nk), C.byref(tp),
                                                   C.byref(nt), _stream_ptr(dev))
            _lib.check(rc, f"scream_render_forward_sharded stage {i}")
        stage(1)
        ko = kp.value - ws.data_ptr()
        reduce_fn(ws[ko:ko + nk.value * 8].view(torch.int64), dist.ReduceOp.MAX)
        stage(2)
        to = tp.value - ws.data_ptr()
        reduce_fn(ws[to:to + nt.value * 4].view(torch.float32), dist.ReduceOp.SUM)
        stage(3)
    return val, win


SHARD_MIN_PTS = 147456          # a slice this large takes the giant-cloud path (scream_render.cu: FS_MIN_AVG_PTS)


def render_cloud_point_sharded(src_local: torch.Tensor, tgt_local: torch.Tensor, src_offset: int, tgt_offset: int,
                               views: torch.Tensor, rho, w: int, group=None, xyz_local: Optional[torch.Tensor] = None,
                               global_counts: Optional[Tuple[int, int]] = None):
    """One giant pair, points sharded over the ranks of `group` (BASELINE config 5; shard.py has the protocol).  Every rank
    passes its rows and their global offsets (xyz_local: the two slices already back to back in one [n+m,3] buffer, if the
    caller keeps them that way -- spares the concatenation); all ranks get the same (imgs [V,2,w,w], val, winner_global
    int64, zrange [V,2]), bit-identical to the unsharded render.  Two collectives per call: all_reduce(MAX) of the
    (-zmin, zmax) pairs, all_reduce(MAX) of the packed 64-bit keys, which scream_pack_keys writes straight into the
    collective's buffer and scream_unpack_keys turns into the images.
    global_counts = (source points, target points) of the WHOLE cloud, cut with shard.balanced_slice: when every rank's
    slice is large enough for the giant-cloud path (w == 64), two more collectives pool the ranks' pre-filter seeds
    (scream_render_forward_sharded), so that the per-rank filter work shrinks with the slice."""
    import torch.distributed as dist
    dev = src_local.device
    n, m = src_local.shape[0], tgt_local.shape[0]
    V = views.shape[0]
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    lib = _lib.load()
    with torch.no_grad():
        if xyz_local is None:
            xyz_local = torch.cat([src_local.float(), tgt_local.float()], 0).contiguous()
        offs = _pair_offsets(n, m, dev)
        if n + m == 0:
            zr = torch.empty((V, 2), dtype=torch.float32, device=dev)
            zr[:, 0], zr[:, 1] = float("inf"), float("-inf")
        else:
            zr = depth_range(xyz_local, offs, views, max_item_pts=n + m)[0]
        if world > 1:                               # min and max in ONE collective: max(-zmin), max(zmax); negation is exact
            zz = torch.stack([-zr[:, 0], zr[:, 1]], dim=-1).contiguous()
            dist.all_reduce(zz, op=dist.ReduceOp.MAX, group=group)
            zr = torch.stack([-zz[:, 0], zz[:, 1]], dim=-1).contiguous()
        with _on_device(dev):
            keys = _empty((V, 2, w, w), torch.int64, dev)
            if n + m == 0:
                keys.zero_()
            else:
                pooled = False
                if global_counts is not None and world > 1 and w == 64:
                    from .shard import slice_counts
                    smallest = min(a + b for a, b in zip(slice_counts(global_counts[0], world), slice_counts(global_counts[1], world)))
                    pooled = smallest >= SHARD_MIN_PTS        # the same answer on every rank
                if pooled:
                    val, win = _sharded_stages(xyz_local, offs, views, rho, w, zr, dist.get_rank(group), world, group)
                else:
                    val, win, _ = render_batch(xyz_local, offs, views, rho, w, raw=True, need_winner=True, zrange_in=zr[None],
                                               max_item_pts=n + m)
                _lib.check(lib.scream_pack_keys(val.data_ptr(), win.data_ptr(), int(src_offset), int(tgt_offset), V, w,
                                                keys.data_ptr(), _stream_ptr(dev)), "scream_pack_keys")
            if world > 1:
                dist.all_reduce(keys, op=dist.ReduceOp.MAX, group=group)
            val_g = _empty((V, 2, w, w), torch.float32, dev)
            imgs = _empty((V, 2, w, w), torch.float32, dev)
            idx_g = _empty((V, 2, w, w), torch.int64, dev)
            _lib.check(lib.scream_unpack_keys(keys.data_ptr(), V, w, val_g.data_ptr(), imgs.data_ptr(), idx_g.data_ptr(),
                                              _stream_ptr(dev)), "scream_unpack_keys")
    return imgs, val_g, idx_g, zr


class RegistrationRender(nn.Module):
    """B200 drop-in for models/render.py:8-73."""

    def __init__(self, rho, w, view="muti", device="cuda:0"):
        super().__init__()
        self.rho, self.w = rho, w
        dev = torch.device(device)           # the reference hard-codes cuda:0 (render.py:12)
        i, j = np.arange(w * w) // w, np.arange(w * w) % w
        pix_xy = torch.from_numpy(np.concatenate([j.reshape(-1, 1), i.reshape(-1, 1)], axis=1)).float()
        # kept for API compatibility (render.py:15-16); the kernels recompute the centres exactly (common.cuh,
        # pixel_centre), so a caller who REPLACES or edits pix_xy is refused at call time instead of silently ignored
        self.pix_xy = ((pix_xy - w // 2 + 0.5) / (w // 2)).to(dev) if dev.type == "cuda" and torch.cuda.is_available() \
            else (pix_xy - w // 2 + 0.5) / (w // 2)
        self._pix_ok = (id(self.pix_xy), self.pix_xy._version, w)
        if view == "muti":
            self.eulers = [np.array([0, 0, 0]), np.array([0, np.pi / 2, 0]), np.array([0, np.pi, 0]),
                           np.array([0, np.pi * 3 / 2, 0]),
                           np.array([0, 0, np.pi / 2]), np.array([0, 0, np.pi * 3 / 2])]
        else:
            self.eulers = [np.array([0, 0, 0])]
        self._cache = {}

    # -- small device-side caches (views by euler content, offsets by sizes) -------------------
    def _views(self, dev) -> torch.Tensor:
        try:                 # the eulers are read on EVERY call (train_kitti.py:48 replaces them); content key, cheap to build
            key = ("v", dev, b"".join([e.tobytes() for e in self.eulers]), self.eulers[0].dtype.num)
        except AttributeError:
            key = ("v", dev, np.asarray(self.eulers, dtype=np.float64).tobytes(), -1)
        hit = self._cache.get(key)
        if hit is None:
            hit = view_matrices(self.eulers).to(dev)
            self._cache = {k: v for k, v in self._cache.items() if k[0] != "v" or k[1] != dev}
            self._cache[key] = hit
        return hit

    def _offsets(self, n: int, m: int, dev) -> torch.Tensor:
        key = ("o", str(dev), n, m)
        hit = self._cache.get(key)
        if hit is None:
            if len(self._cache) > 64:
                self._cache.clear()
            hit = torch.tensor([0, n, n + m], dtype=torch.int64).to(dev)
            self._cache[key] = hit
        return hit

    def _check_pix_xy(self):
        """render.py:42 reads self.pix_xy on every call; here the grid is implied by w.  A pix_xy that is no longer the
        constructor's tensor (or was written in place, or belongs to another w) must equal the canonical grid of the
        CURRENT w -- anything else has no kernel and raises."""
        key = (id(self.pix_xy), getattr(self.pix_xy, "_version", None), self.w)
        if key == self._pix_ok:
            return
        w = int(self.w)
        p = np.arange(w * w)
        want = (torch.from_numpy(np.stack([p % w, p // w], axis=1)).float() - w // 2 + 0.5) / (w // 2)
        got = self.pix_xy
        if not isinstance(got, torch.Tensor) or tuple(got.shape) != (w * w, 2) or not torch.equal(got.detach().float().cpu(), want):
            raise NotImplementedError("pix_xy was changed to something other than the pixel-centre grid of w=%d "
                                      "(render.py:14-16): the B200 kernels render on that grid only" % w)
        self._pix_ok = key

    def _call(self, src, tgt, views, raw, half=False, prerotated_half=False):
        self._check_pix_xy()
        if src.dim() != 2 or tgt.dim() != 2 or src.shape[1] != 3 or tgt.shape[1] != 3:
            raise ValueError("expected [n,3] and [m,3] point tensors")
        if src.shape[0] == 0 or tgt.shape[0] == 0:
            raise RuntimeError("max(): cannot reduce over an empty cloud (render.py:46,49)")
        n, m = src.shape[0], tgt.shape[0]
        if not half and n + m < PAIR_MAX_PTS and not torch.compiler.is_compiling():
            # one pair, one launch: the clouds stay two tensors (render.py:63's cat is implied by the kernel)
            s32 = src if src.dtype == torch.float32 and src.is_contiguous() else src.float().contiguous()
            t32 = tgt if tgt.dtype == torch.float32 and tgt.is_contiguous() else tgt.float().contiguous()
            flags = (_lib.F_RAW if raw else 0) | (_lib.F_HALF if prerotated_half else 0)
            if torch.is_grad_enabled() and (s32.requires_grad or t32.requires_grad):
                return _PairRender.apply(s32, t32, views, int(self.w), float(self.rho), int(flags))
            return _PairRender.run(s32, t32, views, int(self.w), float(self.rho), int(flags), False)[0]
        xyz = torch.cat([src.float(), tgt.float()], dim=0)         # render.py:63
        if prerotated_half:        # render() of fp16 coordinates: only the depth-weight chain of render.py:31 is fp16
            need = torch.is_grad_enabled() and xyz.requires_grad
            imgs, _, _ = torch.ops.scream.render_fwd(xyz.contiguous(), self._offsets(n, m, xyz.device), None, views,
                                                     int(self.w), float(self.rho), int(_lib.F_RAW | _lib.F_HALF),
                                                     bool(need), n + m, None)
            return imgs[0]
        imgs, _, _ = render_batch(xyz, self._offsets(n, m, xyz.device), views, self.rho, self.w,
                                  raw=raw, max_item_pts=n + m, half=half)
        return imgs[0]

    def render(self, x, src_n):
        """[2,w,w] un-normalised image of already-rotated points x [N,3] (render.py:27-59)."""
        _require_cuda(x)
        eye = self._cache.get(("eye", str(x.device)))
        if eye is None:
            eye = torch.eye(3, dtype=torch.float32, device=x.device).unsqueeze(0).contiguous()
            self._cache[("eye", str(x.device))] = eye
        if x.dtype == torch.bfloat16:
            raise NotImplementedError("bf16 coordinates: the reference's KITTI scripts autocast to fp16 (train_kitti.py:154)")
        return self._call(x[:src_n], x[src_n:], eye, raw=True, prerotated_half=x.dtype == torch.float16)[0]

    def forward(self, src_pred, tgt_pcd):
        """[V,2,w,w] fp32 in [-1,1) (render.py:61-73)."""
        _require_cuda(src_pred, tgt_pcd)
        half = False
        if torch.is_autocast_enabled("cuda"):       # train_kitti.py:154-169: render.py:68 runs as an fp16 matmul
            if torch.get_autocast_dtype("cuda") != torch.float16:
                raise NotImplementedError("autocast to bf16: the reference's KITTI scripts autocast to fp16 (train_kitti.py:154)")
            half = True
        return self._call(src_pred, tgt_pcd, self._views(src_pred.device), raw=False, half=half)

    # no parameters / buffers: keep checkpoints of the enclosing model loadable
    def __getstate__(self):
        st = dict(self.__dict__)
        st["_cache"] = {}
        st["_pix_ok"] = None           # ids do not survive pickling: re-checked on the first call
        return st
