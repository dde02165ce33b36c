import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scream_b200 import synth, _lib
from scream_b200.render import depth_range, make_offsets, render_batch
DEV = "cuda:0"
lib = _lib.load()
n = 300
pts = synth.unit_ball(2 * n, 5)
pts[n + 17, 1] = np.float32(np.nan)
print("host bits", hex(int(pts[n + 17].view(np.uint32)[1])))
xyz = torch.from_numpy(pts).to(DEV)
print("dev bits", hex(int(xyz[n + 17].cpu().numpy().view(np.uint32)[1])), "isnan count", int(torch.isnan(xyz).sum()))
offs = make_offsets([(n, n)]).to(DEV)
views = torch.from_numpy(synth.random_views(3, 3)).to(DEV)
N, V = 2 * n, 3
rot = torch.empty((V, N, 3), device=DEV); pv = torch.empty((V, N), device=DEV); zr = torch.empty((1, V, 2), device=DEV)
nws = _lib.workspace_bytes(1, V, 64, N)
ws = torch.zeros(nws, dtype=torch.uint8, device=DEV)
rc = lib.scream_render_transform(xyz.data_ptr(), offs.data_ptr(), N, N, None, views.data_ptr(), 1, V, rot.data_ptr(), pv.data_ptr(),
                                 zr.data_ptr(), ws.data_ptr(), nws, torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print("rc", rc, "rot bits of the NaN point", [hex(int(x)) for x in rot[:, n + 17].cpu().numpy().view(np.uint32).reshape(-1)])
print("zenc words", [hex(int(x)) for x in ws[:V * 8].cpu().numpy().view(np.uint32)])
print("zr bits", [hex(int(x)) for x in zr.cpu().numpy().view(np.uint32).reshape(-1)])
print("pv nan count", int(torch.isnan(pv).sum()))
