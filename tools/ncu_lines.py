#!/usr/bin/env python
"""Per-CUDA-source-line profile of a kernel from an .ncu-rep (needs -lineinfo + --import-source on):
samples, warp instructions, active threads per instruction, shared wavefronts (ideal / actual) and the top stall reasons.
usage: tools/ncu_lines.py gpurun_out/prof_X.ncu-rep [min_pct]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
min_pct = float(sys.argv[2]) if len(sys.argv) > 2 else 0.7
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur_file, hdr, agg = None, None, {}
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No":
        hdr = r; continue
    if hdr is None or len(r) < len(hdr) or not r[0].isdigit():
        continue
    d = dict(zip(hdr, r))
    if d.get("Address") not in ("-", "", None):   # SASS rows belong to the preceding CUDA line: skip (the CUDA row aggregates)
        continue
    key = (cur_file, int(r[0]), r[1].strip()[:90])
    f = lambda k: float(d.get(k) or 0)
    a = agg.setdefault(key, [0, 0, 0, 0, 0, {}])
    a[0] += f("# Samples"); a[1] += f("Instructions Executed"); a[2] += f("Thread Instructions Executed")
    a[3] += f("L1 Wavefronts Shared"); a[4] += f("L1 Wavefronts Shared Ideal")
    for k in hdr:
        if k.startswith("stall_") and "Not Issued" not in k:
            a[5][k] = a[5].get(k, 0) + f(k)
tot_s = sum(a[0] for a in agg.values()); tot_i = sum(a[1] for a in agg.values())
print(f"total samples {tot_s:.0f}  warp-inst {tot_i:.0f}")
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    if 100 * a[0] / tot_s < min_pct:
        continue
    st = sorted(a[5].items(), key=lambda kv: -kv[1])[:3]
    print(f"{key[0]}:{key[1]:4d} smp {100*a[0]/tot_s:5.1f}% inst {100*a[1]/tot_i:5.1f}% thr {a[2]/max(a[1],1):4.1f} "
          f"wf {a[3]/1e6:6.1f}M/{a[4]/1e6:6.1f}M  {' '.join(f'{k[6:]}={v:.0f}' for k, v in st)} | {key[2]}")
