"""Where the final render of a giant cloud's survivor lists spends its time (forward_culled_kernel, -DFC_TIMELINE build):
far / near pixel counts and task times summed over the batches and tiles of a CTA.
    python tools/diag_giant_timeline.py <points> <w> <rho> [views]"""
import ctypes as C, os, subprocess, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from scream_b200.csrc import build as B
lib_path = os.path.join(ROOT, "gpurun_out", "libscream_render_timeline.so")
os.makedirs(os.path.dirname(lib_path), exist_ok=True)
subprocess.check_call([B.find_nvcc()] + B.NVCC_FLAGS + ["-DFC_TIMELINE"] + os.environ.get("SCREAM_DEFS", "").split() + ["-o", lib_path] + B.SOURCES, cwd=B.HERE)
B.LIB = lib_path
from scream_b200 import _lib, synth, render as R
import bench
_lib._can_rebuild = lambda: False
lib = _lib.load()
lib.scream_debug_fc_marks.argtypes = [C.c_void_p]
dev = torch.device("cuda", 0)
n, w, rho = int(sys.argv[1]), int(sys.argv[2]), float(sys.argv[3]); V = int(sys.argv[4]) if len(sys.argv) > 4 else 64
raw = os.environ.get("RAW", "1") == "1"
views = torch.from_numpy(synth.random_views(V, 5)).to(dev)
xyz = torch.cat([bench.counter_ball(0, n // 2, 9000, dev), bench.counter_ball(0, n - n // 2, 9500, dev)], 0)
offs = R.make_offsets([(n // 2, n - n // 2)]).to(dev)
for _ in range(2):
    R.render_batch(xyz, offs, views, rho, w, raw=raw, need_winner=True, max_item_pts=n)
torch.cuda.synchronize()
marks = np.zeros((1024, 24), dtype=np.int64)
lib.scream_debug_fc_marks(marks.ctypes.data)
m = marks[:V * 2]
us = lambda x: x / 1.9e3
print(f"n {n} w {w} rho {rho} raw {raw}: CTAs {len(m)}, CTA total us mean {us(m[:, 8] - m[:, 0]).mean():.0f} max {us(m[:, 8] - m[:, 0]).max():.0f}")
print("  far pixels per CTA (all batches, tiles) mean", m[:, 9].mean(), " near", m[:, 10].mean(), " near->disc fallback", m[:, 11].mean())
print("  far tasks: warp-time sum per CTA (us) mean", us(m[:, 14]).mean(), " longest task (us) max", us(m[:, 12] >> 20).max())
order = np.argsort(-(m[:, 12] >> 20))[:8]
print("  longest far task per CTA (us, batch, first pixel row, col):", [(int(us(m[i, 12] >> 20)), int((m[i, 12] >> 16) & 15), int((m[i, 12] & 0xFFFF) // 64), int((m[i, 12] & 0xFFFF) % 64)) for i in order])
print("  3x3 chunks: warp-time sum per CTA (us) mean", us(m[:, 15]).mean(), " longest (us) max", us(m[:, 13]).max())
