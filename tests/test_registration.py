"""nearest target + weighted Kabsch (pointnet.py:71-74, utils.py:72-78,138-178) -- the oracle against fixtures
frozen from the reference's own functions (CPU), and the CUDA path against both (GPU)."""
import glob
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR, load_golden
from oracle import registration_oracle as reg

NAMES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "registration_*.npz")))
DEV = "cuda:0"


def test_registration_golden_present():
    assert len(NAMES) >= 2


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_reference_functions(name):
    g = load_golden(name)
    T, dist, ind = reg.estimate_transform(torch.from_numpy(g["src"]), torch.from_numpy(g["pred"]), torch.from_numpy(g["tgt"]),
                                          float(g["s"]), float(g["thr"]))
    assert np.array_equal(dist.numpy(), g["dist"]) and np.array_equal(ind.numpy(), g["ind"]) and np.array_equal(T.numpy(), g["T"])
    Tw = reg.rigid_transform_3d(torch.from_numpy(g["wA"]), torch.from_numpy(g["wB"]), torch.from_numpy(g["w"]), float(g["w_thr"]))
    assert np.array_equal(Tw.numpy(), g["Tw"])
    # the fit recovers the pose the fixture was made with (sanity of the fixture itself)
    assert np.abs(g["T"][:3] - g["pose"]).max() < 0.02


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_nearest_target_bit_exact_and_fit_matches(name):
    from scream_b200.registration import estimate_transform, nearest_target, rigid_transform_3d
    g = load_golden(name)
    src, pred, tgt = (torch.from_numpy(g[k]).to(DEV) for k in ("src", "pred", "tgt"))
    dist, ind = nearest_target(pred, tgt, float(g["s"]))
    assert dist.dtype == torch.float32 and ind.dtype == torch.int64
    assert np.array_equal(dist.cpu().numpy(), g["dist"])                       # bit-exact distances ...
    assert np.array_equal(ind.cpu().numpy(), g["ind"])                         # ... and arg-min (first index on ties)
    T, _, _ = estimate_transform(src, pred, tgt, float(g["s"]), float(g["thr"]))
    assert T.shape == (4, 4) and np.abs(T.cpu().numpy() - g["T"]).max() <= 1e-5     # fp64 sums + Jacobi vs fp32 sums + LAPACK
    Tw = rigid_transform_3d(torch.from_numpy(g["wA"]).to(DEV), torch.from_numpy(g["wB"]).to(DEV),
                            torch.from_numpy(g["w"]).to(DEV), float(g["w_thr"]))
    assert Tw.shape == (2, 4, 4) and np.abs(Tw.cpu().numpy() - g["Tw"]).max() <= 2e-5   # item 1: the det < 0 branch
    assert np.array_equal(Tw.cpu().numpy()[:, 3], np.tile(np.float32([0, 0, 0, 1]), (2, 1)))


@pytest.mark.gpu
def test_nearest_target_ties_tiles_and_large_clouds():
    """Exact ties (duplicated targets, a regular grid), more targets than one tile, against the torch-CPU oracle."""
    from scream_b200 import synth
    from scream_b200.registration import nearest_target
    rng = np.random.default_rng(5)
    gx = np.stack(np.meshgrid(*[np.arange(16, dtype=np.float32)] * 3, indexing="ij"), -1).reshape(-1, 3) / np.float32(8) - 1
    tgt = np.concatenate([gx, gx[::3], synth.unit_ball(5000, 9)]).astype(np.float32)      # 4096 + 1366 + 5000 > 4 tiles
    src = np.concatenate([gx[rng.permutation(len(gx))[:700]] + np.float32(0.0625),           # equidistant from 8 grid nodes
                          gx[:300], synth.unit_ball(3000, 10)]).astype(np.float32)
    for s in (1.0, 0.37):
        want_d, want_i = reg.nearest_target(torch.from_numpy(src), torch.from_numpy(tgt), s)
        d, i = nearest_target(torch.from_numpy(src).to(DEV), torch.from_numpy(tgt).to(DEV), s)
        assert np.array_equal(d.cpu().numpy(), want_d.numpy())
        assert np.array_equal(i.cpu().numpy(), want_i.numpy())
    with pytest.raises(RuntimeError):
        nearest_target(torch.from_numpy(src), torch.from_numpy(tgt), 1.0)                    # CPU tensors: no fallback
