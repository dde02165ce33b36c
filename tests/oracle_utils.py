"""numpy front-ends of the plain-C oracle, shared by the CPU and GPU tests (test infrastructure)."""
import ctypes

import numpy as np


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def c_forward(lib, src, tgt, views, rho, w, pose=None):
    src, tgt, views = (np.ascontiguousarray(a, dtype=np.float32) for a in (src, tgt, views))
    pose = None if pose is None else np.ascontiguousarray(pose, dtype=np.float32)
    V = views.shape[0]
    imgs = np.empty((V, 2, w, w), np.float32); val = np.empty_like(imgs); second = np.empty_like(imgs)
    idx = np.empty((V, 2, w, w), np.int32); zr = np.empty((V, 2), np.float32)
    rc = lib.scream_oracle_forward(_p(src), len(src), _p(tgt), len(tgt), _p(pose), _p(views), V, w, float(rho),
                                   _p(imgs), _p(val), _p(idx), _p(second), _p(zr))
    assert rc == 0
    return dict(imgs=imgs, val=val, idx=idx, second=second, zrange=zr)


def c_transform(lib, src, tgt, views, pose=None):
    src, tgt, views = (np.ascontiguousarray(a, dtype=np.float32) for a in (src, tgt, views))
    pose = None if pose is None else np.ascontiguousarray(pose, dtype=np.float32)
    V, N = views.shape[0], len(src) + len(tgt)
    rot = np.empty((V, N, 3), np.float32); pv = np.empty((V, N), np.float32); zr = np.empty((V, 2), np.float32)
    rc = lib.scream_oracle_transform(_p(src), len(src), _p(tgt), len(tgt), _p(pose), _p(views), V, _p(rot), _p(pv), _p(zr))
    assert rc == 0
    return rot, pv, zr


def c_backward(lib, src, tgt, views, rho, w, grad, winner, zrange, pose=None):
    src, tgt, views, grad, zrange = (np.ascontiguousarray(a, dtype=np.float32) for a in (src, tgt, views, grad, zrange))
    winner = np.ascontiguousarray(winner, dtype=np.int32)
    pose = None if pose is None else np.ascontiguousarray(pose, dtype=np.float32)
    gs = np.empty((len(src), 3), np.float64); gt = np.empty((len(tgt), 3), np.float64)
    rc = lib.scream_oracle_backward(_p(src), len(src), _p(tgt), len(tgt), _p(pose), _p(views), views.shape[0], w,
                                    float(rho), _p(grad), _p(winner), _p(zrange), _p(gs), _p(gt))
    assert rc == 0
    return gs, gt


def check_images(got_val, got_idx, want_val, want_idx, want_second, pair_value, rtol=1e-6, atol=1e-37):
    """Parity bar of SURVEY 8c.  Values: <= rtol relative (+ a denormal floor).  Winners: equal to the
    oracle's arg-max except where the oracle's top-2 gap is within a few ulp (near-ties); there the chosen
    point's oracle value must be within that gap of the best.  pair_value(sel) -> oracle values of the
    chosen (pixel, point) pairs for a boolean mask `sel`."""
    err = np.abs(got_val.astype(np.float64) - want_val.astype(np.float64))
    tol = rtol * np.abs(want_val.astype(np.float64)) + atol
    assert (err <= tol).all(), f"value mismatch: max rel {np.max(err / (np.abs(want_val) + 1e-30)):.3e}"
    diff = got_idx != want_idx
    if diff.any():
        chosen = pair_value(diff)
        best = want_val[diff].astype(np.float64)
        gap = np.abs(best - chosen.astype(np.float64))
        assert (gap <= 4 * rtol * np.abs(best) + atol).all(), \
            f"{diff.sum()} winner mismatches, worst rel gap {np.max(gap / (np.abs(best) + 1e-30)):.3e}"
    return int(diff.sum())
