"""CPU suite: the C-ABI library builds, loads and exports every symbol include/scream_render.h declares;
argument validation works without touching a GPU; the drop-in module keeps the reference's surface."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT
from scream_b200 import _lib
from scream_b200.render import RegistrationRender, make_offsets, view_matrices


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "scream_render.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(scream_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported():
    lib = _lib.load()
    names = declared_symbols()
    assert set(names) == set(_lib.EXPORTS)
    for n in names:
        assert getattr(lib, n) is not None
    assert lib.scream_abi_version() == _lib.ABI_VERSION


def test_workspace_query_and_argument_errors():
    lib = _lib.load()
    assert _lib.workspace_bytes(4, 6, 64, 80000) >= 4 * 6 * 2 * 4
    out = ctypes.c_size_t()
    assert lib.scream_render_workspace_bytes(0, 6, 64, 10, ctypes.byref(out)) == -1
    # bad w / NULL pointers are rejected before any CUDA call (so this runs without a GPU)
    one = ctypes.c_void_p(16)
    rc = lib.scream_render_forward(one, one, 10, 10, None, one, 1, 1, 96, 24.0, 0, one, None, None, one, 1 << 20, None)
    assert rc == -1 and b"multiple of 64" in lib.scream_last_error()
    rc = lib.scream_render_forward(None, one, 10, 10, None, one, 1, 1, 64, 24.0, 0, one, None, None, one, 1 << 20, None)
    assert rc == -1
    rc = lib.scream_render_forward(one, one, 10, 10, None, one, 1, 1, 64, 24.0, 0, one, None, None, one, 0, None)
    assert rc == -2
    with pytest.raises(_lib.ScreamError):
        _lib.check(rc, "x")


def test_giant_cloud_path_is_chosen_by_size_and_image_side():
    """Host logic (no GPU): the workspace query tells which calls take the giant-cloud pre-filter -- w = 64 from 147,456
    points per item, w = 128 (four lattice passes) from 2^20, other image sides never; the same sizes decide in the forward call."""
    small = _lib.workspace_bytes(1, 8, 64, 147455)
    assert small < 4096                                          # the encoded depth ranges only
    assert _lib.workspace_bytes(1, 8, 64, 147456) > 100 * small  # seed images, bound tables, survivor lists
    assert _lib.workspace_bytes(2, 8, 64, 2 * 147456 - 1) == _lib.workspace_bytes(2, 8, 64, 1000)   # the AVERAGE item decides
    assert _lib.workspace_bytes(1, 8, 128, (1 << 20) - 1) == small
    big128 = _lib.workspace_bytes(1, 8, 128, 1 << 20)
    assert big128 > 100 * small
    assert big128 == _lib.workspace_bytes(1, 8, 64, 1 << 20)     # the passes reuse one set of buffers: no union lists, no bit maps
    assert _lib.workspace_bytes(1, 8, 192, 1 << 24) == small and _lib.workspace_bytes(1, 8, 256, 1 << 24) == small


def test_module_surface_matches_reference():
    r = RegistrationRender(24, 64, device="cpu")
    assert r.state_dict() == {} and list(r.parameters()) == [] and list(r.buffers()) == []
    assert r.rho == 24 and r.w == 64 and len(r.eulers) == 6 and r.pix_xy.shape == (4096, 2)
    # pix_xy: x from the column, y from the row, no flip (render.py:14-16)
    assert r.pix_xy[0].tolist() == [(-32 + 0.5) / 32, (-32 + 0.5) / 32]
    assert r.pix_xy[65].tolist() == [(1 - 32 + 0.5) / 32, (1 - 32 + 0.5) / 32]
    assert len(RegistrationRender(24, 64, "single", device="cpu").eulers) == 1
    with pytest.raises(RuntimeError, match="CUDA only"):
        r(torch.zeros(20, 3), torch.zeros(20, 3))          # no CPU fallback


def test_view_matrices_are_the_reference_conversion():
    from scipy.spatial.transform import Rotation
    e = [np.array([0, np.pi, 0]), np.array([0.1, 0.2, 0.3])]
    v = view_matrices(e)
    for k in range(2):
        assert torch.equal(v[k], torch.Tensor(Rotation.from_euler("zyx", e[k]).as_matrix()))


def test_make_offsets():
    assert make_offsets([(3, 4), (5, 6)]).tolist() == [0, 3, 7, 12, 18]


def test_custom_ops_have_fake_kernels():
    """scream::render_fwd / render_bwd stay registered for torch.compile (eager calls go through a plain
    autograd.Function around the same host functions): their fake kernels give the shapes and dtypes of the real ones."""
    import torch
    import scream_b200.render  # noqa: F401  (registers the ops)
    xyz = torch.empty(100, 3, device="meta")
    offs = torch.empty(5, dtype=torch.int64, device="meta")
    views = torch.empty(6, 3, 3, device="meta")
    imgs, win, zr = torch.ops.scream.render_fwd(xyz, offs, None, views, 64, 24.0, 0, True, 0, None)
    assert imgs.shape == (2, 6, 2, 64, 64) and win.shape == imgs.shape and win.dtype == torch.int32 and zr.shape == (2, 6, 2)
    _, no_win, _ = torch.ops.scream.render_fwd(xyz, offs, None, views, 64, 24.0, 0, False, 0, None)
    assert no_win.numel() == 0
    g = torch.ops.scream.render_bwd(xyz, offs, None, views, 64, 24.0, 0, imgs, win, zr, 0)
    assert g.shape == xyz.shape
