#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel."""
import collections, csv, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
tot = collections.defaultdict(float); cnt = collections.Counter()
for row in csv.DictReader(lines):
    if row.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    name = re.sub(r'\(.*', '', row['Kernel Name'])[:64]
    v = float(row['Metric Value'].replace(',', '')); u = row['Metric Unit']
    v = v / 1e6 if u in ('nsecond', 'ns') else v / 1e3 if u in ('usecond', 'us') else v
    tot[name] += v; cnt[name] += 1
s = sum(tot.values())
print(f'total {s:.3f} ms, {sum(cnt.values())} launches')
for k, v in sorted(tot.items(), key=lambda x: -x[1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 14]:
    print(f'{v:9.3f} ms {100 * v / s:5.1f}% n={cnt[k]:5d} avg {1e3 * v / cnt[k]:8.1f} us  {k}')
