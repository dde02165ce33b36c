"""Diagnostics (GPU): records per cell of a w = 128 lattice pass's final survivor list (image: view 0, source cloud).
    python tools/surv_cells_w128.py <n_points> [views]"""
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
w, rho = 128, 36.0
nws = _lib.workspace_bytes(1, V, w, n); ws = torch.zeros(nws, dtype=torch.uint8, device=dev)
imgs = torch.empty((1, V, 2, w, w), device=dev); win = torch.empty((1, V, 2, w, w), dtype=torch.int32, device=dev); zr = torch.empty((1, V, 2), device=dev)
rc = lib.scream_render_forward(xyz.data_ptr(), offs.data_ptr(), n, n, None, views.data_ptr(), 1, V, w, rho, 1, imgs.data_ptr(), win.data_ptr(), zr.data_ptr(), ws.data_ptr(), nws, torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize(); assert rc == 0
I = V * 2
cnt_off = al(I * 2 * 4) + al(I * 4096 * 16)
cnt = ws[cnt_off:cnt_off + I * 4].view(torch.int32).cpu().numpy().astype(np.int64) & 0xFFFFFFFF
cnt2 = ws[cnt_off + I * 4:cnt_off + I * 8].view(torch.int32).cpu().numpy().astype(np.int64) & 0xFFFFFFFF
S1 = min(max(49152, n // 16), 1 << 26)
thr_off = cnt_off + al(I * 8 + 64 + 48); surv_off = thr_off + al(I * 4096 * 45 * 4); surv2_off = surv_off + al(I * S1 * 4)
r2 = 0 < cnt2[0] <= (1 << 18)
nn, off = (int(cnt2[0]), surv2_off) if r2 else (int(cnt[0]), surv_off)
ids = ws[off:off + nn * 4].view(torch.int32).long()
p = xyz[:n // 2][ids] @ views[0].T
cj = ((p[:, 0] + 1) * 32).clamp(0, 63).long(); ci = ((p[:, 1] + 1) * 32).clamp(0, 63).long()
h = torch.bincount(ci * 64 + cj, minlength=4096).cpu().numpy()
print(f"list of image 0 (round {'2' if r2 else '1'}): {nn} records; occupied cells {int((h > 0).sum())}; records per cell: max {h.max()}, 99th pct {np.percentile(h[h > 0], 99):.0f}, median {np.median(h[h > 0]):.0f}")
top = np.argsort(-h)[:12]
print("fullest cells (row, col, records):", [(int(t // 64), int(t % 64), int(h[t])) for t in top])
rad = (p[:, 0] ** 2 + p[:, 1] ** 2).sqrt() * 32
print("records beyond radius 31.5 cells:", int((rad > 31.5).sum()), " beyond 31.9:", int((rad > 31.9).sum()), " |x| or |y| > 0.99:", int(((p[:, 0].abs() > 0.99) | (p[:, 1].abs() > 0.99)).sum()))
