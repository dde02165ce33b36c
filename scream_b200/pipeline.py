"""Host-buffer front end: the renderer's share of an adversarial training step, fed from pinned HOST memory.

The reference moves every pair host -> device one at a time (`train_3d_match.py:168-170`, batch 1) and
renders it twice per iteration (`train_3d_match.py:186-201`: "fake" with gradients, "real" with the GT
pose).  `StepPipeline` does the same work for a whole batch of pairs, cut into chunks that travel through
`n_streams` CUDA streams, so the host->device copy of chunk i+1, the kernels of chunk i and the
device->host copy of chunk i-1 overlap (two copy engines + compute).  It calls the same C-ABI entry
points as `render_batch` (the same host functions the torch custom ops wrap) -- there is no other compute path.

    pipe = StepPipeline(sizes, views, rho, w, device)          # sizes: [(n_src, n_tgt)] * B
    pipe.run(h_xyz, h_pose, h_grad, out_fake, out_real, out_gxyz)   # all pinned host tensors
    pipe.synchronize()
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from .render import _bwd_impl, _fwd_impl, make_offsets, render_pair


class StepPipeline:
    def __init__(self, sizes: Sequence[Tuple[int, int]], views: torch.Tensor, rho, w: int, device,
                 chunk_pairs: int = 37, n_streams: int = 6, dense: bool = False, dual: bool = False):
        if not torch.cuda.is_available():
            raise RuntimeError("scream_b200 renders on CUDA only (no CPU fallback)")
        _lib.load()
        self.dev = torch.device(device)
        self.views = views.to(self.dev).contiguous()
        self.rho, self.w = float(rho), int(w)
        self.flags = _lib.F_DENSE if dense else 0
        if dual and dense:
            raise ValueError("dual=True (SCREAM_F_DUAL) is a mode of the culled kernel, not of the dense referee")
        self.dual = bool(dual)                                # fake + real render in one launch
        self.B = len(sizes)
        sizes = [(int(a), int(b)) for a, b in sizes]
        self.chunks: List[Tuple[int, int, int, int, torch.Tensor, int]] = []   # (b0, b1, p0, p1, offs_dev, max_item)
        pts = np.concatenate([[0], np.cumsum([a + b for a, b in sizes])]).astype(np.int64)
        for b0 in range(0, self.B, chunk_pairs):
            b1 = min(self.B, b0 + chunk_pairs)
            offs = make_offsets(sizes[b0:b1]).to(self.dev)
            self.chunks.append((b0, b1, int(pts[b0]), int(pts[b1]), offs, max(a + b for a, b in sizes[b0:b1])))
        self.streams = [torch.cuda.Stream(self.dev) for _ in range(max(1, n_streams))]
        # zrange_init, zrange, forward (x2, or once for a dual launch) + backward scatter
        self.launches_per_run = (4 if self.dual else 7) * len(self.chunks)

    def run(self, h_xyz: torch.Tensor, h_pose: Optional[torch.Tensor], h_grad: torch.Tensor,
            out_fake: torch.Tensor, out_real: torch.Tensor, out_gxyz: torch.Tensor) -> None:
        """Enqueue one step.  h_* are pinned host tensors: xyz [total,3], pose [B,3,4] (GT [R|t] of the
        "real" render), grad [B,V,2,w,w] (upstream gradient of the fake images).  Results land in the
        pinned out_* tensors (fake/real images [B,V,2,w,w], grad_xyz [total,3]) after synchronize()."""
        cur = torch.cuda.current_stream(self.dev)
        for i, (b0, b1, p0, p1, offs, max_item) in enumerate(self.chunks):
            st = self.streams[i % len(self.streams)]
            st.wait_stream(cur)
            with torch.cuda.stream(st):
                x = h_xyz[p0:p1].to(self.dev, non_blocking=True)
                g = h_grad[b0:b1].to(self.dev, non_blocking=True)
                p = h_pose[b0:b1].to(self.dev, non_blocking=True) if h_pose is not None else None
                if self.dual:
                    with torch.no_grad():
                        fake, real, win, zr = render_pair(x, offs, self.views, self.rho, self.w, p, need_winner=True,
                                                          max_item_pts=max_item)
                    out_fake[b0:b1].copy_(fake, non_blocking=True)
                    out_real[b0:b1].copy_(real, non_blocking=True)
                    gx = _bwd_impl(x, offs, None, self.views, self.w, self.rho, self.flags, g, win[0], zr[0], max_item)
                    out_gxyz[p0:p1].copy_(gx, non_blocking=True)
                    continue
                fake, win, zr = _fwd_impl(x, offs, None, self.views, self.w, self.rho, self.flags, True, max_item, None)
                out_fake[b0:b1].copy_(fake, non_blocking=True)
                gx = _bwd_impl(x, offs, None, self.views, self.w, self.rho, self.flags, g, win, zr, max_item)
                out_gxyz[p0:p1].copy_(gx, non_blocking=True)
                real, _, _ = _fwd_impl(x, offs, p, self.views, self.w, self.rho, self.flags, False, max_item, None)
                out_real[b0:b1].copy_(real, non_blocking=True)

    def join(self) -> None:
        """Make the current stream wait for everything enqueued by run() (no host synchronisation)."""
        cur = torch.cuda.current_stream(self.dev)
        for st in self.streams:
            cur.wait_stream(st)

    def synchronize(self) -> None:
        self.join()
        torch.cuda.current_stream(self.dev).synchronize()
