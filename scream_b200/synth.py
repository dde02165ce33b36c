"""Synthetic point clouds of the shapes BASELINE.json names.

The reference's real clouds are not in its tree (download stubs only), so every
benchmark and parity case uses these generators.  Ranges follow the reference's
dataset normalisation: unit ball for 3DMatch (datasets/three_d_match.py:183-190),
bounding-box normalised slab for KITTI (datasets/kitti.py:268-273), xy in [-1,1]
with small z for OpenGG patches (datasets/open_gf.py:8,25).
All generators are numpy-seeded and return contiguous fp32 [n,3] arrays.
"""
from __future__ import annotations

import numpy as np
from scipy.spatial.transform import Rotation


def unit_ball(n: int, seed: int) -> np.ndarray:
    """n points uniform in the unit ball."""
    rng = np.random.default_rng(seed)
    d = rng.standard_normal((n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = rng.random((n, 1)) ** (1.0 / 3.0)
    return np.ascontiguousarray((d * r).astype(np.float32))


def room_shell(n: int, seed: int, noise: float = 0.003) -> np.ndarray:
    """n points on the faces of a random box inside the unit ball (an indoor 'room shell')."""
    rng = np.random.default_rng(seed)
    half = rng.uniform(0.25, 0.55, size=3)
    face = rng.integers(0, 6, size=n)
    p = rng.uniform(-1.0, 1.0, size=(n, 3)) * half
    axis, sign = face // 2, (face % 2) * 2 - 1
    p[np.arange(n), axis] = sign * half[axis]
    p += rng.standard_normal((n, 3)) * noise
    rot = Rotation.random(random_state=seed).as_matrix()
    return np.ascontiguousarray((p @ rot.T).astype(np.float32))


def kitti_slab(n: int, seed: int) -> np.ndarray:
    """Bird's-eye slab: xy uniform in the unit disk, z ~ N(0, 0.02), 5 % of points in raised boxes."""
    rng = np.random.default_rng(seed)
    ang = rng.uniform(0, 2 * np.pi, size=n)
    rad = np.sqrt(rng.random(n))
    p = np.stack([rad * np.cos(ang), rad * np.sin(ang), rng.standard_normal(n) * 0.02], axis=1)
    k = n // 20
    if k:
        centre = rng.uniform(-0.6, 0.6, size=(8, 2))
        which = rng.integers(0, 8, size=k)
        p[:k, :2] = centre[which] + rng.uniform(-0.05, 0.05, size=(k, 2))
        p[:k, 2] = rng.uniform(0.0, 0.15, size=k)
    return np.ascontiguousarray(p.astype(np.float32))


def opengg_patch(side: int, seed: int, coarse: int = 6):
    """(dsm [side*side,3] regular 1 m grid with relief, dem_coarse [coarse*coarse,3]) -- exact ties."""
    rng = np.random.default_rng(seed)
    ax = np.linspace(-1.0, 1.0, side, endpoint=False) + 1.0 / side
    gx, gy = np.meshgrid(ax, ax, indexing="xy")
    z = 0.1 * np.sin(3 * gx) * np.cos(2 * gy) + rng.standard_normal(gx.shape) * 0.01
    dsm = np.stack([gx.ravel(), gy.ravel(), z.ravel()], axis=1)
    cx = np.linspace(-1.0, 1.0, coarse, endpoint=False) + 1.0 / coarse
    hx, hy = np.meshgrid(cx, cx, indexing="xy")
    hz = 0.1 * np.sin(3 * hx) * np.cos(2 * hy) - 0.02
    dem = np.stack([hx.ravel(), hy.ravel(), hz.ravel()], axis=1)
    return (np.ascontiguousarray(dsm.astype(np.float32)),
            np.ascontiguousarray(dem.astype(np.float32)))


def random_pose(seed: int, max_trans: float = 0.15) -> np.ndarray:
    """[3,4] fp32 rigid transform [R|t] (stand-in for the GT poses of 3DMatch val_info.pkl)."""
    rng = np.random.default_rng(seed)
    rot = Rotation.random(random_state=seed).as_matrix()
    t = rng.uniform(-max_trans, max_trans, size=(3, 1))
    return np.ascontiguousarray(np.concatenate([rot, t], axis=1).astype(np.float32))


def random_views(v: int, seed: int) -> np.ndarray:
    """[v,3,3] fp32 random view rotations (cfg5: 64 views)."""
    return np.ascontiguousarray(Rotation.random(v, random_state=seed).as_matrix().astype(np.float32))


def cloud_from_recipe(recipe: str):
    """(src, tgt) from a seeded recipe string: "kind:n:seed|kind:m:seed" with kind in ball / shell / slab, or
    "grid:side:seed" (OpenGG patch: DSM grid + coarse DEM).  The BASELINE-size golden fixtures (tests/golden/full_*.npz)
    store their outputs and this string instead of the clouds."""
    def one(tok):
        kind, n, seed = tok.split(":")
        return {"ball": unit_ball, "shell": room_shell, "slab": kitti_slab}[kind](int(n), int(seed))
    if recipe.startswith("grid:"):
        _, side, seed = recipe.split(":")
        return opengg_patch(int(side), int(seed))
    a, b = recipe.split("|")
    return one(a), one(b)
