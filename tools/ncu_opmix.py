#!/usr/bin/env python
"""Executed-instruction mix by opcode of the first kernel in an ncu report:
python tools/ncu_opmix.py file.ncu-rep [top_n]"""
import collections, csv, io, subprocess, sys
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
iI = hdr.index('Instructions Executed'); iSrc = hdr.index('Source'); iS = hdr.index('# Samples')
mix = collections.Counter(); smp = collections.Counter()
for r in data:
    toks = r[iSrc].strip().split()
    op = toks[1] if toks[0].startswith('@') else toks[0]
    op = op.rstrip(';')
    mix[op] += int(r[iI]); smp[op] += int(r[iS])
tot = sum(mix.values()); ts = sum(smp.values())
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
print(f'{tot} warp-instructions, {ts} samples')
for op, c in mix.most_common(n):
    print(f'{op:40s} {c:10d} {100.0 * c / tot:5.1f}%   samples {100.0 * smp[op] / max(ts, 1):5.1f}%')
