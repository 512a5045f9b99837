#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv): python tools/launch_phases.py file.csv [first-call-only]
Durations are cold-cache and serialised (ncu), so they show the SHARE of each kernel, not the in-situ time."""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    lines = [l for l in open(path) if l.startswith('"')]
    r = csv.reader(lines)
    hdr = next(r)
    ki, vi, gi = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Grid Size')
    rows = [(re.sub(r'\(.*', '', row[ki]), row[gi], float(row[vi].replace(',', '')) / 1e3) for row in r]
    agg = collections.OrderedDict()
    tot = 0.0
    for n, g, t in rows:
        n = re.sub(r'^void ', '', n)
        n = re.sub(r'gq::(<unnamed>::)?', '', n)[:58]
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += t
        tot += t
    print(f'{path}: {len(rows)} launches, {tot / 1e3:.3f} ms serialised')
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f'  {k:60s} {n:5d} {t:10.1f} us  avg {t / n:8.2f}  {100 * t / tot:5.1f} %')


if __name__ == '__main__':
    main()
