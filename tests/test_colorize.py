"""Depth-image colouriser (utils.py:81-91 + the view strip of train_3d_match.py:212-213).
CPU: the numpy restatement against the fixture frozen from the executed reference function.
GPU (-m gpu): the CUDA kernel, through the C ABI, bit-exact against fixture and oracle."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import colorize_oracle as col


def test_oracle_matches_reference_fixture():
    g = load_golden("colorize_strip")
    assert np.array_equal(col.views_to_strip(g["imgs"]), g["strip"])
    assert g["strip"].shape == (64, 8 * 64, 3) and g["strip"].dtype == np.uint8


def test_oracle_tie_and_background_rules():
    x = np.full((1, 2, 64, 64), -1.0, np.float32)          # background in both channels -> black
    assert not col.views_to_strip(x).any()
    x[0, :, 0, 0] = 0.5                                     # tie -> first channel's colour (0, 0.651, 0.929)
    px = col.views_to_strip(x)[0, 0]
    assert px.tolist() == [0, int(np.float32(np.float32(0.75) * np.float32(0.651)) * np.float32(255)),
                           int(np.float32(np.float32(0.75) * np.float32(0.929)) * np.float32(255))]


@pytest.mark.gpu
def test_kernel_matches_fixture_bitwise():
    from scream_b200.colorize import deep_to_img, views_to_strip
    g = load_golden("colorize_strip")
    imgs = torch.from_numpy(g["imgs"]).cuda()
    strip = views_to_strip(imgs)
    assert strip.dtype == torch.uint8 and tuple(strip.shape) == g["strip"].shape
    assert np.array_equal(strip.cpu().numpy(), g["strip"])
    # the drop-in signature: one view, channels last
    one = deep_to_img(imgs[3].permute(1, 2, 0))
    assert np.array_equal(one, g["strip"][:, 3 * 64:4 * 64])


@pytest.mark.gpu
def test_kernel_matches_oracle_on_fresh_batches():
    from scream_b200.colorize import views_to_strip
    rng = np.random.default_rng(9)
    x = rng.uniform(-1, 1, (3, 6, 2, 128, 128)).astype(np.float32)
    x[:, :, 1, ::2] = x[:, :, 0, ::2]                       # ties on every other row
    got = views_to_strip(torch.from_numpy(x).cuda()).cpu().numpy()
    for b in range(3):
        assert np.array_equal(got[b], col.views_to_strip(x[b]))


@pytest.mark.gpu
def test_rendered_views_to_ppm(tmp_path):
    from scream_b200 import synth
    from scream_b200.colorize import views_to_strip, write_ppm
    from scream_b200.render import RegistrationRender
    r = RegistrationRender(24, 64, device="cuda:0")
    imgs = r(torch.from_numpy(synth.unit_ball(800, 1)).cuda(), torch.from_numpy(synth.room_shell(700, 2)).cuda())
    strip = views_to_strip(imgs)
    assert np.array_equal(strip.cpu().numpy(), col.views_to_strip(imgs.cpu().numpy()))
    path = tmp_path / "views.ppm"
    write_ppm(str(path), strip)
    raw = open(path, "rb").read()
    head = b"P6\n384 64\n255\n"
    assert raw.startswith(head) and len(raw) == len(head) + 64 * 384 * 3
