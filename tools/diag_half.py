"""Determinism stress of the fp16 (autocast) mode: the module under autocast against the oracle, with allocator churn (NaN-filled\nblocks freed between calls).  ITERS=2 under compute-sanitizer (initcheck / racecheck)."""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import render_oracle as orc
from scream_b200 import synth, _lib
from scream_b200.render import RegistrationRender, make_offsets, render_batch, view_matrices
DEV="cuda:0"
src, tgt = synth.unit_ball(600, 1), synth.room_shell(500, 2)
views_t = view_matrices(RegistrationRender(24, 64, device="cpu").eulers)
want = orc.forward_instrumented(torch.from_numpy(src), torch.from_numpy(tgt), 24, 64, views_t, half=True)
r = RegistrationRender(24, 64, device=DEV)
s = torch.from_numpy(src).cuda().requires_grad_(True); t = torch.from_numpy(tgt).cuda().requires_grad_(True)
bad = 0
for it in range(int(os.environ.get("ITERS", "300"))):
    if it % 3 == 0:
        imgs = r(s, t); imgs.backward(torch.randn_like(imgs)); del imgs
        junk = []
        for k in range(24):                      # freed blocks of many size classes, filled with NaN / ones / random bits / zeros
            nby = int(2 ** np.random.uniform(9, 24))
            j = torch.empty((nby // 4,), dtype=torch.int32, device=DEV)
            mode = (it // 3 + k) % 4
            if mode == 0: j.fill_(0x7fc00000)
            elif mode == 1: j.fill_(-1)
            elif mode == 2: j.random_(-2**31, 2**31 - 1)
            else: j.zero_()
            junk.append(j)
        del junk
    with torch.autocast("cuda", dtype=torch.float16), torch.no_grad():
        h = r(s.detach(), t.detach())
    e = (h.cpu() - want["imgs"]).abs()
    if e.max() > 2e-6:
        bad += 1
        b = (e > 2e-6).nonzero()
        if bad <= 5: print("iter", it, "mismatches", len(b), b[:4].tolist(), e.max().item())
print("bad iterations", bad, "of 300")
