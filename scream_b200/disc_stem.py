"""The discriminator's first layer on the renderer's output: models/gan.py:39,
`nn.Conv2d(input_nc, ndf, kernel_size=4, stride=2, padding=1)` + `nn.LeakyReLU(0.2, True)`, which loss.py:50-57 applies to
the rendered views (`g_loss = -mean(D(imgs))`).  CUDA kernels behind the C ABI (`scream_disc_stem_forward / _backward`),
differentiable w.r.t. the images, the weight and the bias.  The renderer emits 2 channels (render.py:58-59) -- the shipped
discriminator is built with input_nc=3 (loss.py:19-24) and cannot take them; build it with input_nc=2.

    stem = DiscStem.from_discriminator(D)            # D = NLayerDiscriminator(input_nc=2, ...): shares D.main[0]'s parameters
    feat = stem(imgs)                                # == D.main[1](D.main[0](imgs)); then D.main[2:](feat)
"""
from __future__ import annotations

import torch
from torch import nn

from . import _lib
from .render import _empty, _on_device, _ptr, _stream_ptr


class _DiscStemFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        if not x.is_cuda:
            raise RuntimeError("scream_b200 runs on CUDA only (no CPU fallback): got a %s tensor" % x.device)
        lead = x.shape[:-3]
        x4 = x.reshape((-1,) + tuple(x.shape[-3:])).float().contiguous()
        w = weight.float().contiguous()
        b = None if bias is None else bias.float().contiguous()
        N, C, H, W = x4.shape
        F = w.shape[0]
        if tuple(w.shape) != (F, C, 4, 4):
            raise ValueError(f"weight must be [F,{C},4,4] (gan.py:38-39: kernel 4), got {tuple(w.shape)}")
        dev = x4.device
        with _on_device(dev):
            y = _empty((N, F, H // 2, W // 2), torch.float32, dev)
            rc = _lib.load().scream_disc_stem_forward(x4.data_ptr(), w.data_ptr(), _ptr(b), N, C, F, H, W, y.data_ptr(), _stream_ptr(dev))
        _lib.check(rc, "scream_disc_stem_forward")
        ctx.save_for_backward(x4, w, y)
        ctx.meta = (tuple(x.shape), bias is not None)
        return y.view(tuple(lead) + (F, H // 2, W // 2))

    @staticmethod
    def backward(ctx, dy):
        x4, w, y = ctx.saved_tensors
        xshape, has_bias = ctx.meta
        N, C, H, W = x4.shape
        F = w.shape[0]
        dy = dy.reshape(y.shape).float().contiguous()
        dev = x4.device
        need_x, need_w, need_b = ctx.needs_input_grad[0], ctx.needs_input_grad[1], has_bias and ctx.needs_input_grad[2]
        with _on_device(dev):
            dx = _empty(tuple(x4.shape), torch.float32, dev) if need_x else None
            dw = _empty(tuple(w.shape), torch.float32, dev) if (need_w or need_b) else None
            db = _empty((F,), torch.float32, dev) if need_b else None
            rc = _lib.load().scream_disc_stem_backward(x4.data_ptr(), w.data_ptr(), y.data_ptr(), dy.data_ptr(), N, C, F, H, W,
                                                       _ptr(dx), _ptr(dw), _ptr(db), _stream_ptr(dev))
        _lib.check(rc, "scream_disc_stem_backward")
        return (None if dx is None else dx.view(xshape)), (dw if need_w else None), db


def disc_stem(imgs: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor = None) -> torch.Tensor:
    """leaky_relu(conv2d(imgs, weight, bias, stride=2, padding=1), 0.2) -- gan.py:39.  imgs [...,C,H,W] (the renderer's
    [V,2,w,w] or [B,V,2,w,w]); weight [F,C,4,4]; returns [...,F,H/2,W/2]."""
    return _DiscStemFn.apply(imgs, weight, bias)


class DiscStem(nn.Module):
    """gan.py:39 as one module.  Parameters named like nn.Conv2d's (`weight`, `bias`), initialised like the reference
    (weights_init, gan.py:6-9: N(0, 0.02) weights; Conv2d's default bias)."""

    def __init__(self, input_nc: int = 2, ndf: int = 64):
        super().__init__()
        conv = nn.Conv2d(input_nc, ndf, kernel_size=4, stride=2, padding=1)
        nn.init.normal_(conv.weight.data, 0.0, 0.02)
        self.weight, self.bias = conv.weight, conv.bias

    @classmethod
    def from_discriminator(cls, disc: nn.Module) -> "DiscStem":
        conv = disc.main[0]
        m = cls.__new__(cls)
        nn.Module.__init__(m)
        m.weight, m.bias = conv.weight, conv.bias              # shared, not copied: the optimiser of D keeps working
        return m

    def forward(self, imgs: torch.Tensor) -> torch.Tensor:
        return disc_stem(imgs, self.weight, self.bias)
