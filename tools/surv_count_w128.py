"""Diagnostics (GPU): survivor counts of the w = 128 lattice passes (the last pass's two rounds) beside w = 64.
    python tools/surv_count_w128.py <n_points> [views]"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from scream_b200 import _lib, synth
from scream_b200.render import make_offsets
n = int(sys.argv[1]); V = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device('cuda:0')
xyz = torch.cat([bench.counter_ball(0, n // 2, 9000, dev), bench.counter_ball(0, n - n // 2, 9500, dev)], 0)
offs = make_offsets([(n // 2, n - n // 2)]).to(dev); views = torch.from_numpy(synth.random_views(V, 5)).to(dev)
lib = _lib.load()
al = lambda x: (x + 255) // 256 * 256
for w, rho in ((64, 24.0), (128, 36.0), (128, 24.0)):
    nws = _lib.workspace_bytes(1, V, w, n); ws = torch.zeros(nws, dtype=torch.uint8, device=dev)
    imgs = torch.empty((1, V, 2, w, w), device=dev); win = torch.empty((1, V, 2, w, w), dtype=torch.int32, device=dev); zr = torch.empty((1, V, 2), device=dev)
    rc = lib.scream_render_forward(xyz.data_ptr(), offs.data_ptr(), n, n, None, views.data_ptr(), 1, V, w, rho, 1, imgs.data_ptr(), win.data_ptr(), zr.data_ptr(), ws.data_ptr(), nws, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize(); assert rc == 0
    I = V * 2
    cnt_off = al(I * 2 * 4) + al(I * 4096 * 16)
    cnt = ws[cnt_off:cnt_off + I * 4].view(torch.int32).cpu().numpy().astype(np.int64) & 0xFFFFFFFF
    cnt2 = ws[cnt_off + I * 4:cnt_off + I * 8].view(torch.int32).cpu().numpy().astype(np.int64) & 0xFFFFFFFF
    print(f"w {w} rho {rho}: (last pass) round-1 mean {cnt.mean():.0f} max {cnt.max()}  round-2 mean {cnt2.mean():.0f} max {cnt2.max()}")
    wn = win[0].cpu().numpy()
    print("   distinct winners per image (mean)", np.mean([len(np.unique(wn[v, c])) for v in range(V) for c in range(2)]))
    del ws, imgs, win
