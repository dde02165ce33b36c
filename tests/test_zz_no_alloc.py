"""scream_render_forward_noalloc (include/scream_render.h): the forward render with its scratch in a static device array.

No counterpart in the reference.  Called through the C ABI with ctypes, on two
calls in a row with different inputs, and compared with scream_render_forward on a caller workspace: the same kernels run on
the same inputs, so the tolerance is zero -- images, winner maps and depth ranges must be bit-identical.  The file sorts last
on purpose.  The entry point itself allocates nothing; the test's own tensors come from torch.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from scream_b200 import _lib, synth
from scream_b200.render import make_offsets, view_matrices

pytestmark = pytest.mark.gpu

EULERS = [[0.0, 0.0, 0.0], [0.0, 90.0, 0.0], [90.0, 0.0, 0.0], [0.0, 180.0, 0.0], [0.0, 270.0, 0.0], [270.0, 0.0, 0.0]]


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _batch(seed, sizes):
    clouds = []
    for k, (n, m) in enumerate(sizes):
        clouds.append((synth.unit_ball(n, seed + 2 * k), synth.room_shell(m, seed + 2 * k + 1)))
    xyz = torch.from_numpy(np.concatenate([np.concatenate(p) for p in clouds]).astype(np.float32)).cuda()
    offs = make_offsets([(len(s), len(t)) for s, t in clouds]).cuda()
    return xyz, offs


def _render(entry, xyz, offs, views, B, V, w, rho, flags, pose=None, ws=None):
    lib = _lib.load()
    dev = xyz.device
    imgs = torch.full((B, V, 2, w, w), float("nan"), device=dev)
    win = torch.full((B, V, 2, w, w), -7, dtype=torch.int32, device=dev)
    zr = torch.full((B, V, 2), float("nan"), device=dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    total = int(xyz.shape[0])
    head = (_ptr(xyz), _ptr(offs), total, 0, _ptr(pose), _ptr(views), B, V, w, float(rho), flags, _ptr(imgs), _ptr(win), _ptr(zr))
    if entry == "noalloc":
        rc = lib.scream_render_forward_noalloc(*head, stream)
    else:
        rc = lib.scream_render_forward(*head, _ptr(ws), ws.numel(), stream)
    _lib.check(rc, entry)
    torch.cuda.synchronize()
    return imgs.cpu(), win.cpu(), zr.cpu()


def _same(a, b):
    return all(np.array_equal(x.numpy().view(np.uint32) if x.dtype == torch.float32 else x.numpy(),
                              y.numpy().view(np.uint32) if y.dtype == torch.float32 else y.numpy()) for x, y in zip(a, b))


def test_noalloc_forward_equals_workspace_forward_on_two_calls_in_a_row():
    lib = _lib.load()
    views = view_matrices(EULERS).cuda().contiguous()
    V = views.shape[0]
    # call 1: three ragged pairs, w = 64, with a pose; call 2: other clouds, other batch size, w = 128, RAW values.
    # The second call finds the first call's ranges in the static array: every word it reads must be rewritten first.
    cases = [(11, [(700, 500), (64, 1300), (2048, 16)], 64, 24.0, 0, True),
             (29, [(1500, 900), (333, 777)], 128, 36.0, _lib.F_RAW, False)]
    for seed, sizes, w, rho, flags, posed in cases:
        xyz, offs = _batch(seed, sizes)
        B = len(sizes)
        pose = None
        if posed:
            g = torch.Generator().manual_seed(seed)
            q, _ = torch.linalg.qr(torch.randn(B, 3, 3, generator=g))
            pose = torch.cat([q, 0.05 * torch.randn(B, 3, 1, generator=g)], dim=2).float().cuda().contiguous()
        ws = torch.empty(_lib.workspace_bytes(B, V, w, int(xyz.shape[0])), dtype=torch.uint8, device="cuda")
        want = _render("workspace", xyz, offs, views, B, V, w, rho, flags, pose, ws)
        got = _render("noalloc", xyz, offs, views, B, V, w, rho, flags, pose)
        assert _same(got, want), f"seed {seed}: scream_render_forward_noalloc differs from scream_render_forward"
        assert torch.isfinite(got[0]).all() and int(got[1].min()) >= 0        # outputs were written, not left at the fill
        again = _render("noalloc", xyz, offs, views, B, V, w, rho, flags, pose)
        assert _same(again, want), f"seed {seed}: the second call in a row differs"

    # beyond the static array: refused with a clear error, nothing launched
    xyz, offs = _batch(5, [(40, 40)])
    Vbig = _lib.NOALLOC_WS_BYTES // 16 + 1
    many = views[:1].repeat(Vbig, 1, 1).contiguous()
    imgs = torch.zeros(1, device="cuda")
    rc = lib.scream_render_forward_noalloc(_ptr(xyz), _ptr(offs), 80, 0, None, _ptr(many), 1, Vbig, 64, 24.0, 0,
                                           _ptr(imgs), None, None, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -4 and b"static array" in lib.scream_last_error()
    torch.cuda.synchronize()
