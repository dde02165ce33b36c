"""CPU oracle of the step that follows the generator -- TEST INFRASTRUCTURE ONLY (see render_oracle.py).

torch-CPU restatement of utils.py:72-78 (square_distance) and utils.py:138-178 (rigid_transform_3d, with integrate_trans,
utils.py:118-135) with the same ATen ops in the same order; oracle/make_golden.py --only-registration executes the
reference's own functions (their source segments, compiled unmodified: `import utils` needs open3d) and asserts equality
before freezing tests/golden/registration_*.npz.
"""
from __future__ import annotations

import torch


def square_distance(src: torch.Tensor, dst: torch.Tensor) -> torch.Tensor:
    """[B,N,M] = -2 src.dst^T + |src|^2 + |dst|^2, utils.py:72-78 (in-place adds in this order)."""
    B, N, _ = src.shape
    M = dst.shape[1]
    d = -2 * torch.matmul(src, dst.permute(0, 2, 1))
    d += torch.sum(src ** 2, -1).view(B, N, 1)
    d += torch.sum(dst ** 2, -1).view(B, 1, M)
    return d


def nearest_target(src_pred: torch.Tensor, tgt: torch.Tensor, s: float):
    """pointnet.py:71."""
    return square_distance(src_pred[None] / s, tgt[None] / s)[0].min(dim=1)


def rigid_transform_3d(A: torch.Tensor, B: torch.Tensor, weights=None, weight_threshold=0) -> torch.Tensor:
    """utils.py:138-178 -> [bs,4,4]."""
    bs = A.shape[0]
    w = torch.ones_like(A[:, :, 0]) if weights is None else weights.clone()
    w[w < weight_threshold] = 0
    ca = torch.sum(A * w[:, :, None], dim=1, keepdim=True) / (torch.sum(w, dim=1, keepdim=True)[:, :, None] + 1e-6)
    cb = torch.sum(B * w[:, :, None], dim=1, keepdim=True) / (torch.sum(w, dim=1, keepdim=True)[:, :, None] + 1e-6)
    H = (A - ca).permute(0, 2, 1) @ torch.diag_embed(w) @ (B - cb)
    U, S, V = torch.svd(H)
    eye = torch.eye(3)[None].repeat(bs, 1, 1)
    eye[:, -1, -1] = torch.det(V @ U.permute(0, 2, 1))
    R = V @ eye @ U.permute(0, 2, 1)
    t = cb.permute(0, 2, 1) - R @ ca.permute(0, 2, 1)
    T = torch.eye(4)[None].repeat(bs, 1, 1)                 # integrate_trans, utils.py:118-135
    T[:, :3, :3] = R
    T[:, :3, 3:4] = t.view(-1, 3, 1)
    return T


def estimate_transform(src, src_pred, tgt, s, dist_threshold):
    """pointnet.py:71-74."""
    dist, ind = nearest_target(src_pred, tgt, s)
    valid = dist < dist_threshold
    return rigid_transform_3d(src[None][:, valid], tgt[None][:, ind[valid]])[0], dist, ind
