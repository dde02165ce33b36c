"""Depth images -> colour strip on the device: the reference's `deep_to_img` (utils.py:81-91) and the
`np.concatenate([...], axis=1)` around it (train_3d_match.py:212-213, train_kitti.py:211-212,
visualize_depth_image.py:27-28), as one CUDA kernel behind the C ABI (`scream_depth_to_rgb8`).

The reference then calls cv2.imwrite / cv2.imshow; the viewer window is unavailable on this box, so the
headless product is the uint8 buffer (and, if asked, a binary PPM/PNG-free file written from the host copy).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib


def views_to_strip(imgs: torch.Tensor) -> torch.Tensor:
    """imgs [V,2,w,w] or [B,V,2,w,w] fp32 cuda (the renderer's normalised output) -> uint8 [w, V*w, 3]
    (or [B, w, V*w, 3]) on the same device, equal to
    np.concatenate([deep_to_img(imgs[j].permute(1, 2, 0)) for j in range(V)], axis=1)."""
    if not imgs.is_cuda:
        raise RuntimeError("scream_b200 colourises on CUDA only (no CPU fallback): got a %s tensor" % imgs.device)
    single = imgs.dim() == 4
    x = (imgs[None] if single else imgs).detach().float().contiguous()
    if x.dim() != 5 or x.shape[2] != 2 or x.shape[3] != x.shape[4]:
        raise ValueError("expected [V,2,w,w] or [B,V,2,w,w]")
    B, V, _, w, _ = x.shape
    lib = _lib.load()
    with torch.cuda.device(x.device):
        out = torch.empty((B, w, V * w, 3), dtype=torch.uint8, device=x.device)
        rc = lib.scream_depth_to_rgb8(x.data_ptr(), B, V, w, out.data_ptr(), torch.cuda.current_stream(x.device).cuda_stream)
    _lib.check(rc, "scream_depth_to_rgb8")
    return out[0] if single else out


def deep_to_img(deep_img: torch.Tensor) -> np.ndarray:
    """Drop-in for utils.py:81-91: deep_img [w,w,2] (one view, channels last) -> numpy uint8 [w,w,3]."""
    if deep_img.dim() != 3 or deep_img.shape[2] != 2:
        raise ValueError("expected a [w,w,2] tensor")
    return views_to_strip(deep_img.permute(2, 0, 1)[None]).cpu().numpy()


def write_ppm(path: str, strip) -> None:
    """Headless stand-in for cv2.imwrite(path, img_cv): binary PPM (P6), components in the array's order."""
    a = strip.cpu().numpy() if isinstance(strip, torch.Tensor) else np.asarray(strip)
    if a.ndim != 3 or a.shape[2] != 3 or a.dtype != np.uint8:
        raise ValueError("expected uint8 [h,w,3]")
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (a.shape[1], a.shape[0]))
        f.write(np.ascontiguousarray(a).tobytes())
