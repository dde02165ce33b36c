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


def check_images(got_val, got_idx, want, pair_value=None, rtol=1e-6, atol=1e-37, budget=None, label=""):
    """Parity bar of SURVEY 8c / north_star, winner maps included.

    want: dict with the oracle's val, idx and either `second` (runner-up VALUE per pixel) or `tie` (bool: second == val);
    optional near_pix / near_idx (oracle/render_oracle.py: for every pixel whose top-2 gap is within 2 ulp, the other
    points that reach the runner-up value).
      values   <= rtol relative (+ a denormal floor)
      winners  got_idx == want_idx wherever the oracle's two best values are EXACTLY equal (torch.max: lowest index)
               or more than 2 ulp apart; in between (a near-tie: exp() is an opaque <= 1 ulp routine on both sides) the
               chosen point must be one of the oracle's listed runner-ups -- or, without lists, its oracle value
               (pair_value(mask) -> values of the chosen (pixel, point) pairs) must equal the oracle's runner-up value.
    Returns the number of near-tie pixels that took the runner-up; `budget` bounds it (0: winner maps bit-exact)."""
    want_val, want_idx = np.asarray(want["val"]), np.asarray(want["idx"]).astype(np.int64)
    got_idx = np.asarray(got_idx).astype(np.int64)
    err = np.abs(got_val.astype(np.float64) - want_val.astype(np.float64))
    tol = rtol * np.abs(want_val.astype(np.float64)) + atol
    assert (err <= tol).all(), f"{label}: value mismatch: max rel {np.max(err / (np.abs(want_val) + 1e-30)):.3e}"
    diff = got_idx != want_idx
    n_diff = int(diff.sum())
    if n_diff:
        if "second" in want:
            gap = want_val.astype(np.float64) - np.asarray(want["second"]).astype(np.float64)
            tie = gap == 0
            near = (gap > 0) & (gap <= 2 * np.spacing(np.abs(want_val)).astype(np.float64))
        else:
            tie = np.asarray(want["tie"]).astype(bool)
            near = np.zeros_like(tie)
            if "near_pix" in want:
                near.reshape(-1)[np.asarray(want["near_pix"]).astype(np.int64)] = True
        assert not (diff & tie).any(), f"{label}: {int((diff & tie).sum())} exact ties did not resolve to the lowest index"
        assert not (diff & ~near).any(), \
            f"{label}: {int((diff & ~near).sum())} winners differ at pixels whose oracle top-2 gap exceeds 2 ulp"
        if "near_pix" in want:
            lists = {int(p): set(int(i) for i in row if i >= 0) for p, row in zip(np.asarray(want["near_pix"]), np.asarray(want["near_idx"]))}
            for p in np.flatnonzero(diff.reshape(-1)):
                assert int(got_idx.reshape(-1)[p]) in lists.get(int(p), ()), \
                    f"{label}: pixel {p}: chose {int(got_idx.reshape(-1)[p])}, oracle runner-ups {sorted(lists.get(int(p), ()))}"
        else:
            assert pair_value is not None and "second" in want
            chosen = pair_value(diff)
            assert np.array_equal(chosen, np.asarray(want["second"])[diff]), f"{label}: a chosen point is not the oracle's runner-up"
    if budget is not None:
        assert n_diff <= budget, f"{label}: {n_diff} near-tie winners differ from the oracle, budget {budget}"
    print(f"[parity] {label}: {n_diff} near-tie winner(s) took the oracle's runner-up, of {diff.size} pixels")
    return n_diff
