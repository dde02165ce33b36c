"""Phase timeline of forward_culled_kernel for one pair per call (cfg1): a diagnostic build of the library (-DFC_TIMELINE)
whose CTAs record the SM cycle counter at the phase boundaries.  usage: python tools/diag_cfg1_timeline.py [views]"""
import ctypes as C, os, subprocess, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from scream_b200.csrc import build as B
lib_path = os.path.join(ROOT, "gpurun_out", "libscream_render_timeline.so")
os.makedirs(os.path.dirname(lib_path), exist_ok=True)
subprocess.check_call([B.find_nvcc()] + B.NVCC_FLAGS + ["-DFC_TIMELINE"] + os.environ.get("SCREAM_DEFS", "").split() + ["-o", lib_path] + B.SOURCES, cwd=B.HERE)
B.LIB = lib_path                      # the loader below takes the diagnostic build
from scream_b200 import _lib, synth, render as R
_lib._can_rebuild = lambda: False
lib = _lib.load()
lib.scream_debug_fc_marks.argtypes = [C.c_void_p]
dev = torch.device("cuda", 0)
V = int(sys.argv[1]) if len(sys.argv) > 1 else 6
src = torch.from_numpy(synth.unit_ball(10000, 1)).to(dev); tgt = torch.from_numpy(synth.unit_ball(10000, 2)).to(dev)
r = R.RegistrationRender(24, 64, device=dev)
if V != 6:
    r.eulers = r.eulers[:V]
names = ["start", "zrange", "P1", "P2", "P3", "P3b", "classify", "worklist", "near"]
BATCH = int(sys.argv[2]) if len(sys.argv) > 2 else 0       # > 0: that many pairs through render_batch (one CTA per image, cfg2's mode)
if BATCH:
    xyz = torch.cat([torch.cat([torch.from_numpy(synth.unit_ball(10000, 2 * i + 1)), torch.from_numpy(synth.unit_ball(10000, 2 * i + 2))]) for i in range(BATCH)]).to(dev)
    offs = R.make_offsets([(10000, 10000)] * BATCH).to(dev)
    vm = R.view_matrices(r.eulers).to(dev)
    class _B:
        def __call__(self, a, b):
            return R.render_batch(xyz, offs, vm, 24, 64, need_winner=True)[0]
    r = _B()
for it in range(4):
    with torch.no_grad():
        out = r(src, tgt)
    torch.cuda.synchronize()
marks = np.zeros((1024, 24), dtype=np.int64)
lib.scream_debug_fc_marks(marks.ctypes.data)
n = min(1024, V * 2 * BATCH) if BATCH else V * 2 * min(64, 148 // (V * 2), int(os.environ.get("SCREAM_MAX_PARTS", "64")))
m = marks[:n]
d = np.diff(m[:, :9], axis=1) / 1.9e3          # us at ~1.9 GHz
import time
def bench_calls():
    for _ in range(20):
        with torch.no_grad(): r(src, tgt)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(200):
        with torch.no_grad(): r(src, tgt)
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / 200 * 1e6
print("defs", os.environ.get("SCREAM_DEFS", ""), " us per call (instrumented build, wall):", round(bench_calls(), 1))
print("CTAs", n, "phase durations (us): mean / max over CTAs")
for k in range(8):
    print(f"  {names[k + 1]:9s} {d[:, k].mean():7.2f} {d[:, k].max():7.2f}")
tot = (m[:, 8] - m[:, 0]) / 1.9e3
print(f"  total     {tot.mean():7.2f} {tot.max():7.2f}")
print("far pixels per CTA mean/max", m[:, 9].mean(), m[:, 9].max(), " near", m[:, 10].mean(), m[:, 10].max(), " fallback", m[:, 11].mean(), m[:, 11].max())
print("far batch us: max over CTAs of max", m[:, 12].max() / 1.9e3, "mean of max", m[:, 12].mean() / 1.9e3, " mean task", (m[:, 14].sum() / np.maximum(1, np.ceil(m[:, 9] / np.maximum(1, np.minimum(32, (m[:, 9] + 31) // 32))).sum())) / 1.9e3)
print("3x3 chunk us: max over CTAs of max", m[:, 13].max() / 1.9e3, "mean of max", m[:, 13].mean() / 1.9e3, " mean task", m[:, 15].sum() / (128 * V * 2 * max(1, BATCH)) / 1.9e3)
if m[:, 21].max() > 0:
    print("cluster sort (us, mean): P1 work", ((m[:, 21] - m[:, 1]) / 1.9e3).mean(), "P1 barrier", ((m[:, 2] - m[:, 21]) / 1.9e3).mean(),
          "P3 work", ((m[:, 22] - m[:, 3]) / 1.9e3).mean(), "P3 barrier", ((m[:, 23] - m[:, 22]) / 1.9e3).mean(), "copy", ((m[:, 4] - m[:, 23]) / 1.9e3).mean())
ns = np.maximum(1, m[:, 19].sum())
print("warp search per pixel (us): guess", m[:, 16].sum() / ns / 1.9e3, "rows", m[:, 17].sum() / ns / 1.9e3, "merge", m[:, 18].sum() / ns / 1.9e3, " searches per CTA", m[:, 19].mean(), " mean reach (rows)", m[:, 20].sum() / ns)
worst = np.argsort(-d[:, 7])[:6]
for cta in worst:
    print(f"  cta {cta}: near phase {d[cta, 7]:.2f} us, near {m[cta, 10]}, fallback {m[cta, 11]}, worklist {d[cta, 6]:.2f}, far {m[cta, 9]}")
