import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scream_b200 import synth, render as R, _lib
dev = torch.device("cuda", 0)
src = torch.from_numpy(synth.unit_ball(10000, 1)).to(dev); tgt = torch.from_numpy(synth.unit_ball(10000, 2)).to(dev)
r = R.RegistrationRender(24, 64, device=dev)
g = torch.randn((6, 2, 64, 64), device=dev)
def fb():
    s = src.detach().requires_grad_(True)
    r(s, tgt).backward(g)
def bench(tag):
    for _ in range(50): fb()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(300): fb()
    torch.cuda.synchronize(); print(tag, round((time.perf_counter() - t0) / 300 * 1e6, 1), "us")
bench("current")
# variant: time pieces on the main thread
s = src.detach().requires_grad_(True)
t0 = time.perf_counter()
for _ in range(300): out = r(s, tgt)
torch.cuda.synchronize(); print("forward only (grad mode)", round((time.perf_counter() - t0) / 300 * 1e6, 1))
out = r(s, tgt)
t0 = time.perf_counter()
for _ in range(300): out.backward(g, retain_graph=True)
torch.cuda.synchronize(); print("backward only", round((time.perf_counter() - t0) / 300 * 1e6, 1))
orig = R._PairRender.backward
import cProfile, pstats
def prof_bwd(ctx, gi):
    pr.enable(); o = orig(ctx, gi); pr.disable(); return o
pr = cProfile.Profile()
R._PairRender.backward = staticmethod(prof_bwd)
for _ in range(300): out.backward(g, retain_graph=True)
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(10)
