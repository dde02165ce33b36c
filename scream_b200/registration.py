"""The step that follows the generator: nearest target + weighted Kabsch on the device.

Mirrors the reference's helpers (same names and argument meaning):
  square_distance(...)[0].min(dim=1)  utils.py:72-78 + pointnet.py:71   -> nearest_target(src_pred, tgt, s)
  rigid_transform_3d(A, B, weights, weight_threshold)  utils.py:138-178 -> rigid_transform_3d (no CPU SVD, no host sync)
  pointnet.py:71-74 (train_kitti.py:171-175 with its own threshold)     -> estimate_transform (both, no boolean-mask sync)
C ABI: scream_nearest_target, scream_rigid_transform (include/scream_render.h).  CUDA only; no CPU fallback.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib
from .render import _ptr, _require_cuda, _stream_ptr


def nearest_target(src_pred: torch.Tensor, tgt: torch.Tensor, s: float = 1.0) -> Tuple[torch.Tensor, torch.Tensor]:
    """(dist [n] fp32, ind [n] int64) == square_distance(src_pred[None] / s, tgt[None] / s)[0].min(dim=1), bit for bit
    (utils.py:72-78, pointnet.py:71), without the n x m matrix."""
    _require_cuda(src_pred, tgt)
    if src_pred.dim() != 2 or tgt.dim() != 2 or src_pred.shape[1] != 3 or tgt.shape[1] != 3:
        raise ValueError("expected [n,3] and [m,3] point tensors")
    if src_pred.shape[0] == 0 or tgt.shape[0] == 0:
        raise RuntimeError("min(): cannot reduce over an empty cloud (pointnet.py:71)")
    a, b = src_pred.detach().float().contiguous(), tgt.detach().float().contiguous()
    n, m, dev = a.shape[0], b.shape[0], a.device
    with torch.cuda.device(dev):
        dist = torch.empty((n,), dtype=torch.float32, device=dev)
        ind = torch.empty((n,), dtype=torch.int64, device=dev)
        ws = torch.empty((n,), dtype=torch.int64, device=dev)
        rc = _lib.load().scream_nearest_target(_ptr(a), n, _ptr(b), m, float(s), _ptr(dist), _ptr(ind), _ptr(ws), n * 8,
                                               _stream_ptr(dev))
    _lib.check(rc, "scream_nearest_target")
    return dist, ind


def _fit(A, B, b_index, weights, valid_dist, valid_thr, weight_threshold):
    _require_cuda(A, B, b_index, weights, valid_dist)
    A, B = A.detach().float().contiguous(), B.detach().float().contiguous()
    if A.dim() != 3 or B.dim() != 3 or A.shape[2] != 3 or B.shape[2] != 3 or A.shape[0] != B.shape[0]:
        raise ValueError("expected A [bs,k,3] and B [bs,k',3]")
    bs, k, dev = A.shape[0], A.shape[1], A.device
    if b_index is None and B.shape[1] != k:
        raise ValueError("A and B need the same number of correspondences")
    if k == 0:
        raise RuntimeError("rigid_transform_3d: no correspondences")
    if weights is not None:
        weights = weights.detach().float().contiguous()
    if b_index is not None:
        b_index = b_index.to(torch.int64).contiguous()
    if valid_dist is not None:
        valid_dist = valid_dist.detach().float().contiguous()
    with torch.cuda.device(dev):
        T = torch.empty((bs, 4, 4), dtype=torch.float32, device=dev)
        ws = torch.empty((bs * 16,), dtype=torch.float64, device=dev)
        rc = _lib.load().scream_rigid_transform(_ptr(A), _ptr(B), _ptr(b_index), _ptr(weights), _ptr(valid_dist),
                                                float(valid_thr), float(weight_threshold), bs, k, B.shape[1],
                                                _ptr(T), _ptr(ws), bs * 128, _stream_ptr(dev))
    _lib.check(rc, "scream_rigid_transform")
    return T


def rigid_transform_3d(A: torch.Tensor, B: torch.Tensor, weights: Optional[torch.Tensor] = None,
                       weight_threshold: float = 0) -> torch.Tensor:
    """[bs,4,4] transforms taking A [bs,k,3] onto B [bs,k,3] (utils.py:138-178; weights [bs,k], those below
    weight_threshold count 0).  Unlike the reference it does not write the clipped weights back into `weights`."""
    return _fit(A, B, None, weights, None, 0.0, weight_threshold)


def estimate_transform(src: torch.Tensor, src_pred: torch.Tensor, tgt: torch.Tensor, s: float = 1.0,
                       dist_threshold: float = 0.075) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """pointnet.py:71-74 for one pair: nearest target of every predicted point, correspondences closer than
    dist_threshold (squared, in units of s), Kabsch of src [n,3] onto the matched targets.  Returns
    (T [4,4], dist [n], ind [n]); nothing touches the host."""
    dist, ind = nearest_target(src_pred, tgt, s)
    T = _fit(src[None], tgt[None], ind[None], None, dist[None], dist_threshold, 0.0)
    return T[0], dist, ind
