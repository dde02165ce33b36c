"""The discriminator's first layer on the renderer's output (gan.py:39: Conv2d(C, 64, 4, 2, 1) + LeakyReLU(0.2)).
CPU: the torch restatement against the fixture frozen from the reference's own NLayerDiscriminator modules.
GPU (-m gpu): the CUDA kernels through the C ABI.  Floating point: cuDNN / oneDNN do not fix their accumulation order, so the
bar is 1e-5 relative to the largest magnitude of the tensor compared (forward and the three gradients)."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import disc_stem_oracle as dso

RTOL = 1e-5


def test_oracle_matches_reference_fixture():
    g = load_golden("disc_stem")
    y = dso.disc_stem(torch.from_numpy(g["x"]), torch.from_numpy(g["weight"]), torch.from_numpy(g["bias"]))
    assert np.array_equal(y.numpy(), g["y"])
    assert g["y"].shape == (2, 64, 32, 32) and (g["y"] < 0).any() and (g["y"] > 0).any()


def _close(got, want, what):
    scale = np.abs(want).max()
    err = np.abs(got - want).max()
    assert err <= RTOL * scale, f"{what}: max err {err:.3e} vs scale {scale:.3e}"


@pytest.mark.gpu
def test_kernels_match_reference_forward_and_autograd():
    from scream_b200.disc_stem import disc_stem
    g = load_golden("disc_stem")
    x = torch.from_numpy(g["x"]).cuda().requires_grad_(True)
    w = torch.from_numpy(g["weight"]).cuda().requires_grad_(True)
    b = torch.from_numpy(g["bias"]).cuda().requires_grad_(True)
    y = disc_stem(x, w, b)
    assert y.shape == g["y"].shape and y.dtype == torch.float32 and y.is_contiguous()
    _close(y.detach().cpu().numpy(), g["y"], "forward")
    y.backward(torch.from_numpy(g["grad_out"]).cuda())
    _close(x.grad.cpu().numpy(), g["grad_x"], "grad x")
    _close(w.grad.cpu().numpy(), g["grad_w"], "grad weight")
    _close(b.grad.cpu().numpy(), g["grad_b"], "grad bias")


@pytest.mark.gpu
def test_on_rendered_batch_against_torch_modules():
    """The renderer's [B,V,2,w,w] output goes in as it is; w = 128; gradients reach the point clouds through both ops."""
    from scream_b200 import synth
    from scream_b200.disc_stem import DiscStem
    from scream_b200.render import RegistrationRender
    torch.manual_seed(3)
    r = RegistrationRender(36, 128, device="cuda:0")
    s = torch.from_numpy(synth.unit_ball(3000, 1)).cuda().requires_grad_(True)
    t = torch.from_numpy(synth.room_shell(2500, 2)).cuda()
    stem = DiscStem(2, 64).cuda()
    imgs = r(s, t)
    feat = stem(imgs)
    assert feat.shape == (6, 64, 64, 64)
    want = dso.disc_stem(imgs.detach().cpu(), stem.weight.detach().cpu(), stem.bias.detach().cpu())
    _close(feat.detach().cpu().numpy(), want.numpy(), "forward w=128")
    (-feat.mean()).backward()                                  # loss.py:57  g_loss = -mean(D(fake))
    assert s.grad is not None and torch.isfinite(s.grad).all() and float(s.grad.abs().max()) > 0
    # the same gradient through torch's own conv on the device
    s2 = s.detach().clone().requires_grad_(True)
    imgs2 = r(s2, t)
    ref = torch.nn.functional.leaky_relu(torch.nn.functional.conv2d(imgs2, stem.weight.detach(), stem.bias.detach(), stride=2, padding=1), 0.2)
    (-ref.mean()).backward()
    _close(s.grad.cpu().numpy(), s2.grad.cpu().numpy(), "grad through the renderer")
    # 5-D input and a discriminator's own parameters
    five = stem(imgs.detach()[None])
    assert five.shape == (1, 6, 64, 64, 64) and torch.equal(five[0], feat.detach())
