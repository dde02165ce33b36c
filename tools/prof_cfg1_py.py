"""Host-side cost of one-pair calls: cProfile over 3000 forward and 1500 forward+backward calls of the drop-in module."""
import cProfile, os, pstats, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scream_b200 import synth
from scream_b200.render import RegistrationRender
dev = torch.device("cuda", 0)
src = torch.from_numpy(synth.unit_ball(10000, 1)).to(dev); tgt = torch.from_numpy(synth.unit_ball(10000, 2)).to(dev)
r = RegistrationRender(24, 64, device=dev)
g = torch.randn((6, 2, 64, 64), device=dev)
def fwd(n):
    with torch.no_grad():
        for _ in range(n):
            r(src, tgt)
    torch.cuda.synchronize()
def fb(n):
    for _ in range(n):
        s = src.detach().requires_grad_(True)
        r(s, tgt).backward(g)
    torch.cuda.synchronize()
fwd(200); fb(200)
for name, fn, n in (("forward", fwd, 3000), ("forward+backward", fb, 1500)):
    pr = cProfile.Profile(); pr.enable(); fn(n); pr.disable()
    print("=====", name, "per call us:", 1e6 * pstats.Stats(pr).total_tt / n)
    pstats.Stats(pr).sort_stats("tottime").print_stats(14)
