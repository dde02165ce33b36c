"""Times the discriminator-stem kernels one by one through the C ABI (512 x 6 rendered views' worth of input).
usage: python tools/bench_disc_stem.py"""
import ctypes as C, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scream_b200 import _lib
lib = _lib.load()
dev = torch.device("cuda", 0)
N, Cc, F, H, W = 3072, 2, 64, 64, 64
torch.manual_seed(0)
x = torch.randn(N, Cc, H, W, device=dev); w = torch.randn(F, Cc, 4, 4, device=dev) * 0.1; b = torch.randn(F, device=dev) * 0.1
y = torch.empty(N, F, H // 2, W // 2, device=dev); dy = torch.randn_like(y)
dx = torch.empty_like(x); dw = torch.empty_like(w); db = torch.empty_like(b)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
fwd = lambda: _lib.check(lib.scream_disc_stem_forward(p(x), p(w), p(b), N, Cc, F, H, W, p(y), st), "fwd")
bwd = lambda a, c, d: (lambda: _lib.check(lib.scream_disc_stem_backward(p(x), p(w), p(y), p(dy), N, Cc, F, H, W, p(a), p(c), p(d), st), "bwd"))
gb = (y.numel() * 4 * 2 + x.numel() * 4) / 1e9
print(f"forward {timed(fwd):.3f} ms; dx only {timed(bwd(dx, None, None)):.3f} ms; dW, db only {timed(bwd(None, dw, db)):.3f} ms; all {timed(bwd(dx, dw, db)):.3f} ms"
      f"  (y + dy + x = {gb:.2f} GB: {gb / 6.5:.3f} ms at the HBM peak per backward kernel)")
