import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scream_b200 import synth
from scream_b200.render import RegistrationRender
dev = torch.device("cuda", 0)
src = torch.from_numpy(synth.unit_ball(10000, 1)).to(dev); tgt = torch.from_numpy(synth.unit_ball(10000, 2)).to(dev)
r = RegistrationRender(24, 64, device=dev)
g = torch.randn((6, 2, 64, 64), device=dev)
for _ in range(30):
    s = src.detach().requires_grad_(True)
    r(s, tgt).backward(g)
torch.cuda.synchronize()
