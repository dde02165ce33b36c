"""N > 1 path on CPU: world_size-2 (and 3) gloo process groups run scream_b200/shard.py's protocols with a
small numpy fp32 renderer standing in for the CUDA kernels (the same callable interface the product
passes).  Checks: item sharding + gather is a pure partition; the point-sharded merge (depth-range
all_reduce, int64 key all_reduce MAX) is BIT-IDENTICAL to the unsharded render, including lowest-index
tie-breaks and all-zero pixels."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from scream_b200 import shard, synth

W, RHO = 64, 24


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _centres(w):
    c = ((np.arange(w, dtype=np.float32) - np.float32(w // 2)) + np.float32(0.5)) / np.float32(w // 2)
    return np.tile(c, w), np.repeat(c, w)


def np_range(pts, views):
    """[V,2] zmin/zmax of fl(fma chain) rotated depth (render.py:30-31, 68)."""
    out = np.empty((len(views), 2), np.float32)
    for v, R in enumerate(views):
        z = (pts.astype(np.float32) @ R[2].astype(np.float32))
        out[v] = (z.min(), z.max()) if len(z) else (np.inf, -np.inf)
    return out


def np_render(src, tgt, views, zr, rho=RHO, w=W):
    """(val [V,2,w,w] f32, winner [V,2,w,w] i32 local index, -1 for an empty cloud): render.py:27-59 with
    an externally supplied depth range."""
    cx, cy = _centres(w)
    V = len(views)
    val = np.zeros((V, 2, w * w), np.float32)
    win = np.full((V, 2, w * w), -1, np.int32)
    k = np.float32(rho * rho)
    for v, R in enumerate(views):
        R = R.astype(np.float32)
        for c, pts in enumerate((src, tgt)):
            if len(pts) == 0:
                continue
            X = pts.astype(np.float32) @ R.T
            pv = np.float32(1) - (X[:, 2] - zr[v, 0]) / np.float32(np.float64(zr[v, 1]) - np.float64(zr[v, 0]))
            d2 = (X[:, 0:1] - cx[None]) ** 2 + (X[:, 1:2] - cy[None]) ** 2
            t = pv[:, None] * np.exp(-d2 / np.float32(2) * k)
            win[v, c] = t.argmax(0)                      # first index on ties, like torch.max(dim=0)
            val[v, c] = t.max(0)
    return val.reshape(V, 2, w, w), win.reshape(V, 2, w, w)


def _clouds():
    # a regular grid (exact ties) + duplicated points (ties across shards) + a far corner (all-zero pixels)
    g, _ = synth.opengg_patch(16, 3)
    src = np.concatenate([g * np.float32(0.4), g[:40] * np.float32(0.4)], 0)
    tgt = synth.unit_ball(301, 5) * np.float32(0.35)
    views = np.stack([np.eye(3, dtype=np.float32), synth.random_views(1, 2)[0]])
    return src.astype(np.float32), tgt.astype(np.float32), views


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        src, tgt, views = _clouds()
        # ---- point sharding ----
        slo, shi = shard.balanced_slice(len(src), rank, world)
        tlo, thi = shard.balanced_slice(len(tgt), rank, world)

        def range_fn(s, t):
            return torch.from_numpy(np_range(np.concatenate([s.numpy(), t.numpy()], 0), views))

        def render_fn(s, t, zr):
            v, wn = np_render(s.numpy(), t.numpy(), views, zr.numpy())
            return torch.from_numpy(v), torch.from_numpy(wn)

        imgs, val, idx, zr = shard.render_cloud_point_sharded(
            torch.from_numpy(src[slo:shi]), torch.from_numpy(tgt[tlo:thi]), slo, tlo, range_fn, render_fn)
        loc = shard.local_winner(idx, slo, shi - slo, tlo, thi - tlo)
        # ---- item sharding ----
        n_items = 5
        ilo, ihi = shard.balanced_slice(n_items, rank, world)
        local = torch.stack([torch.full((2, 3), float(i)) for i in range(ilo, ihi)]) if ihi > ilo else torch.empty((0, 2, 3))
        allimgs = shard.gather_items(local, n_items)
        stats = shard.reduce_stats({"n": ihi - ilo, "sum": float(sum(range(ilo, ihi)))}, "cpu")
        ret[rank] = dict(imgs=imgs.numpy(), val=val.numpy(), idx=idx.numpy(), zr=zr.numpy(), loc=loc.numpy(),
                         allimgs=allimgs.numpy(), stats=stats, span=(slo, shi, tlo, thi))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_point_and_item_sharding_gloo(world):
    port = _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    src, tgt, views = _clouds()
    zr = np_range(np.concatenate([src, tgt], 0), views)
    val, win = np_render(src, tgt, views, zr)
    assert (val == 0).any(), "fixture should contain all-zero pixels"
    for r in range(world):
        out = ret[r]
        assert np.array_equal(out["zr"], zr)
        assert np.array_equal(out["val"].view(np.int32), val.view(np.int32))          # bit-identical values
        assert np.array_equal(out["idx"], win.astype(np.int64))                         # identical winners (ties -> lowest index)
        assert np.array_equal(out["imgs"], (val - np.float32(0.5)) / np.float32(0.5))
        slo, shi, tlo, thi = out["span"]
        for c, (lo, hi) in enumerate(((slo, shi), (tlo, thi))):
            mine = (win[:, c] >= lo) & (win[:, c] < hi)
            assert np.array_equal(out["loc"][:, c] >= 0, mine)
            assert np.array_equal(out["loc"][:, c][mine], (win[:, c] - lo)[mine])
        assert np.array_equal(out["allimgs"], np.stack([np.full((2, 3), float(i), np.float32) for i in range(5)]))
        assert out["stats"] == {"n": 5.0, "sum": 10.0}
    # every pixel's winner is owned by exactly one rank
    owners = sum((ret[r]["loc"] >= 0).astype(int) for r in range(world))
    assert (owners == 1).all()


def test_partition_and_key_order():
    for n in (0, 1, 7, 64, 1000003):
        for world in (1, 2, 3, 8):
            spans = [shard.balanced_slice(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1 and sizes == shard.slice_counts(n, world)
    # key order == (value, then lower index); round trip
    val = torch.tensor([0.0, 1e-45, 0.25, 0.25, 1.0, 0.25])
    idx = torch.tensor([3, 9, 7, 2, 0, -1])
    key = shard.pack_keys(val, idx)
    order = sorted(range(6), key=lambda i: int(key[i]))
    assert order == [5, 0, 1, 2, 3, 4]
    v2, i2 = shard.unpack_keys(key[:5])
    assert torch.equal(v2, val[:5]) and torch.equal(i2, idx[:5].long())
