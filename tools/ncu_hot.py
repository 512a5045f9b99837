#!/usr/bin/env python
"""Hottest SASS instructions of the first kernel in an ncu report (by warp-stall samples):
python tools/ncu_hot.py file.ncu-rep [top_n]"""
import csv, io, subprocess, sys
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv', '--print-source', 'sass'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None; data = []; kernels = 0
for r in rows:
    if r and r[0] == 'Kernel Name':
        kernels += 1
        if kernels > 1: break
        continue
    if r and r[0] == 'Address':
        hdr = r; continue
    if hdr and len(r) == len(hdr):
        data.append(r)
iS = hdr.index('# Samples'); iI = hdr.index('Instructions Executed'); iSrc = hdr.index('Source')
tot = sum(int(r[iS]) for r in data)
print(f'{len(data)} SASS instructions, {tot} samples, {sum(int(r[iI]) for r in data)} warp-instructions executed')
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
for idx, r in sorted(enumerate(data), key=lambda t: -int(t[1][iS]))[:n]:
    print(f'{idx:5d} {int(r[iS]):6d} {100.0 * int(r[iS]) / tot:5.1f}%  exec {int(r[iI]):7d}  {r[iSrc].strip()[:90]}')
