"""fp16 (autocast) input mode, SURVEY 8 f2 (train_kitti.py:154-169): the CUDA path against fixtures frozen from the
reference executed under torch.autocast (CPU, fp16) and against the explicit-rounding oracle (oracle/render_oracle.py)."""
import numpy as np
import pytest
import torch

from conftest import half_golden_names, load_golden
from oracle import render_oracle as orc
from oracle_utils import check_images
from scream_b200 import _lib, synth
from scream_b200.render import RegistrationRender, make_offsets, render_batch, view_matrices

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
NAMES = half_golden_names()


def test_half_golden_present():
    assert len(NAMES) >= 2


@pytest.mark.parametrize("dense", [True, False])
@pytest.mark.parametrize("name", NAMES)
def test_half_forward_matches_reference_under_autocast(name, dense):
    g = load_golden(name)
    w, rho = int(g["w"]), float(g["rho"])
    xyz = torch.from_numpy(np.concatenate([g["src"], g["tgt"]], 0)).to(DEV)
    offs = make_offsets([(len(g["src"]), len(g["tgt"]))]).to(DEV)
    views = torch.from_numpy(g["views"]).to(DEV)
    val, idx, zr = render_batch(xyz, offs, views, rho, w, raw=True, dense=dense, need_winner=True, half=True)
    imgs, _, _ = render_batch(xyz, offs, views, rho, w, dense=dense, need_winner=False, half=True)
    assert val.shape == (1,) + g["val"].shape and idx.shape == val.shape and zr.shape == (1,) + g["zrange"].shape
    assert np.array_equal(zr[0].cpu().numpy(), g["zrange"])                     # bit-exact (fp16-valued) z range
    idx_np = idx[0].cpu().numpy()
    check_images(val[0].cpu().numpy(), idx_np, g, label=f"{name} half dense={dense}")
    assert np.abs(imgs[0].cpu().numpy() - g["imgs"]).max() <= 2e-6


@pytest.mark.parametrize("name", NAMES)
def test_half_module_under_cuda_autocast_forward_and_backward(name):
    """The drop-in module switches to the fp16 arithmetic exactly when the reference would: inside
    torch.autocast("cuda", dtype=float16).  Gradient: the reference's autocast autograd rounds its intermediates to
    fp16; the product passes fp32 gradients straight through the casts -> equal to 2^-11-ish of the largest gradient
    (4.3e-4 measured on the CPU against the same fixture); tolerance 2e-3."""
    g = load_golden(name)
    w, rho = int(g["w"]), float(g["rho"])
    r = RegistrationRender(rho if rho != int(rho) else int(rho), w, device=DEV)
    r.eulers = list(g["eulers"])
    s = torch.from_numpy(g["src"]).to(DEV).requires_grad_(True)
    t = torch.from_numpy(g["tgt"]).to(DEV).requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.float16):
        imgs = r(s, t)
    assert imgs.dtype == torch.float32 and imgs.shape == g["imgs"].shape and imgs.is_contiguous()
    assert np.abs(imgs.detach().cpu().numpy() - g["imgs"]).max() <= 2e-6
    plain = r(s.detach(), t.detach())                                            # outside autocast: the fp32 path
    assert (plain != imgs.detach()).float().mean().item() > 0.2
    imgs.backward(torch.from_numpy(g["grad_out"]).to(DEV))
    scale = max(np.abs(g["grad_src"]).max(), np.abs(g["grad_tgt"]).max())
    assert np.abs(s.grad.cpu().numpy() - g["grad_src"]).max() <= 2e-3 * scale
    assert np.abs(t.grad.cpu().numpy() - g["grad_tgt"]).max() <= 2e-3 * scale
    with torch.autocast("cuda", dtype=torch.bfloat16), pytest.raises(NotImplementedError):
        r(s, t)


def test_half_rotate_bit_exact_and_batch_layout():
    """scream_half_rotate against the oracle's fp16 matmul restatement, items of unequal sizes; the batched render
    equals the items rendered one by one; the culled kernel equals the dense referee bit for bit."""
    sizes = [(700, 900), (1, 2500), (3000, 40)]
    clouds = []
    for i, (n, m) in enumerate(sizes):
        clouds += [synth.unit_ball(n, 300 + i) * np.float32(1.7), synth.kitti_slab(m, 400 + i)]
    xyz_np = np.concatenate(clouds, 0)
    total, B = len(xyz_np), len(sizes)
    views_t = view_matrices(RegistrationRender(24, 64, device="cpu").eulers)
    V = views_t.shape[0]
    xyz, offs, views = torch.from_numpy(xyz_np).to(DEV), make_offsets(sizes).to(DEV), views_t.to(DEV)
    xr = torch.empty((V * total, 3), device=DEV)
    offs_bv = torch.empty((2 * B * V + 1,), dtype=torch.int64, device=DEV)
    rc = _lib.load().scream_half_rotate(xyz.data_ptr(), offs.data_ptr(), total, 0, views.data_ptr(), B, V,
                                        xr.data_ptr(), offs_bv.data_ptr(), torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "scream_half_rotate")
    xr_np, ob, o = xr.cpu().numpy(), offs_bv.cpu().numpy(), make_offsets(sizes).numpy()
    assert ob[0] == 0 and ob[-1] == V * total and (np.diff(ob) >= 0).all()
    for b in range(B):
        pts = torch.from_numpy(xyz_np[o[2 * b]:o[2 * b + 2]])
        for v in range(V):
            j = b * V + v
            assert ob[2 * j + 1] - ob[2 * j] == sizes[b][0] and ob[2 * j + 2] - ob[2 * j + 1] == sizes[b][1]
            assert np.array_equal(xr_np[ob[2 * j]:ob[2 * j + 2]], orc.rotate_half(views_t[v], pts).numpy()), (b, v)
    val, idx, zr = render_batch(xyz, offs, views, 24, 64, raw=True, need_winner=True, half=True)
    vd, idd, zd = render_batch(xyz, offs, views, 24, 64, raw=True, need_winner=True, half=True, dense=True)
    assert torch.equal(val, vd) and torch.equal(idx, idd) and torch.equal(zr, zd)
    for b in range(B):
        one = torch.from_numpy(xyz_np[o[2 * b]:o[2 * b + 2]]).to(DEV)
        v1, i1, z1 = render_batch(one, make_offsets([sizes[b]]).to(DEV), views, 24, 64, raw=True, need_winner=True, half=True)
        assert torch.equal(v1[0], val[b]) and torch.equal(i1[0], idx[b]) and torch.equal(z1[0], zr[b])


def test_half_render_of_fp16_coordinates_and_refusals():
    """render(x, src_n) with fp16 x (what render.py:27 receives under autocast): only the depth-weight chain differs."""
    g = load_golden(NAMES[0])
    w, rho = int(g["w"]), float(g["rho"])
    pts = torch.from_numpy(np.concatenate([g["src"], g["tgt"]], 0))
    x = orc.rotate_half(torch.from_numpy(g["views"][0]), pts)
    want = orc.render_instrumented(x, len(g["src"]), rho, w, half=True)
    r = RegistrationRender(rho if rho != int(rho) else int(rho), w, device=DEV)
    got = r.render(x.half().to(DEV), len(g["src"]))
    assert got.shape == (2, w, w) and got.dtype == torch.float32
    assert np.abs(got.cpu().numpy() - want["val"].numpy()).max() <= 1e-6
    xyz, offs = pts.to(DEV), make_offsets([(len(g["src"]), len(g["tgt"]))]).to(DEV)
    pose = torch.from_numpy(synth.random_pose(3))[None].contiguous().to(DEV)
    with pytest.raises(ValueError):
        render_batch(xyz, offs, torch.from_numpy(g["views"]).to(DEV), rho, w, pose=pose, half=True)
