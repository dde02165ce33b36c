"""fp16 (autocast) mode, cold: the autocast render is the FIRST CUDA work of a fresh process (optionally with
SCREAM_POISON=1: every buffer the path allocates pre-filled with 0xFF), compared with the explicit-rounding oracle.
Prints every mismatching pixel with the winner the kernel chose and the oracle's winner / runner-up.  Exit code 1 on
any mismatch.  usage: [SCREAM_POISON=1] python tools/diag_half_cold.py [seed]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import render_oracle as orc                                  # noqa: E402  (checker)
from scream_b200 import synth                                             # noqa: E402
from scream_b200.render import RegistrationRender, make_offsets, render_batch, view_matrices   # noqa: E402

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
src, tgt = synth.unit_ball(600, seed), synth.room_shell(500, seed + 1)
views_t = view_matrices(RegistrationRender(24, 64, device="cpu").eulers)
want = orc.forward_instrumented(torch.from_numpy(src), torch.from_numpy(tgt), 24, 64, views_t, half=True)
DEV = "cuda:0"
xyz = torch.from_numpy(np.concatenate([src, tgt], 0)).to(DEV)
offs = make_offsets([(600, 500)]).to(DEV)
views = views_t.to(DEV)
bad_total = 0
for it in range(3):
    val, idx, zr = render_batch(xyz, offs, views, 24, 64, raw=True, need_winner=True, half=True, max_item_pts=1100)
    v, i = val[0].cpu(), idx[0].cpu()
    e = (v - want["val"]).abs()
    bad = (e > 1e-6 * want["val"].abs() + 1e-37).nonzero()
    zbad = not torch.equal(zr[0].cpu(), want["zrange"])
    if len(bad) or zbad:
        bad_total += 1
        print(f"call {it}: {len(bad)} value mismatches, zrange differs: {zbad}")
        for b in bad[:8]:
            t = tuple(b.tolist())
            print("   pixel", t, "got", float(v[t]), "idx", int(i[t]), "| oracle", float(want["val"][t]), "idx", int(want["idx"][t]),
                  "second", float(want["second"][t]))
print("cold half diag:", "MISMATCH" if bad_total else "clean", "poison" if os.environ.get("SCREAM_POISON") else "")
sys.exit(1 if bad_total else 0)
