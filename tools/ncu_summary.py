#!/usr/bin/env python
"""Summarise an .ncu-rep: key raw metrics + samples/instructions per barrier-delimited SASS segment.
usage: tools_ncu_summary.py gpurun_out/prof_X.ncu-rep [lo hi]   (lo..hi: dump SASS lines with counts)"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
keys = ['gpu__time_duration.sum', 'sm__cycles_elapsed.max', 'smsp__inst_executed.sum', 'dram__bytes_read.sum ', 'dram__bytes_write.sum ',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__issue_active.avg.pct', 'launch__registers_per_thread ',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ', 'launch__grid_size',
        'sm__warps_active.avg.pct_of_peak_sustained_active']
for h, u, v in zip(hdr, units, vals):
    if any((h + ' ').startswith(k) or h == k.strip() for k in keys):
        print(f"{h} [{u}] = {v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr, data = rows[1], rows[2:]
iS, iI, iT, iA = hdr.index('# Samples'), hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed'), hdr.index('Avg. Threads Executed')
tot = sum(int(r[iS]) for r in data); totI = sum(int(r[iI]) for r in data)
print('total samples', tot, 'warp-inst', totI)
seg = 0; acc = [0, 0, 0]; start = 0
for k, r in enumerate(data):
    acc[0] += int(r[iS]); acc[1] += int(r[iI]); acc[2] += int(r[iT])
    if 'BAR.SYNC' in r[1] or k == len(data) - 1:
        print(f'seg {seg} lines {start}-{k}: samples {acc[0]} ({100*acc[0]/tot:.1f}%) inst {acc[1]} ({100*acc[1]/totI:.1f}%) thr/inst {acc[2]/max(acc[1],1):.1f}')
        seg += 1; acc = [0, 0, 0]; start = k + 1
if len(sys.argv) > 3:
    for k in range(int(sys.argv[2]), int(sys.argv[3])):
        r = data[k]
        print(k, r[1][:64].ljust(64), r[iS], r[iI], r[iA])
