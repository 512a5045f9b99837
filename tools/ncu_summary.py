#!/usr/bin/env python
"""Compact summary of an `ncu --set full` report: python tools/ncu_summary.py file.ncu-rep [more.ncu-rep ...]
One line per captured launch with the numbers bench.py's roofline needs (duration, DRAM traffic, tensor-pipe
activity, occupancy), in plain units.  `--stalls` adds the top warp-stall reasons."""
import csv
import io
import subprocess
import sys

KEYS = {
    'gpu__time_duration.sum': 'dur',
    'dram__bytes_read.sum': 'dram_rd',
    'dram__bytes_write.sum': 'dram_wr',
    'lts__t_bytes.sum': 'l2_bytes',
    'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active': 'tensor_pct',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed': 'dram_pct',
    'sm__warps_active.avg.pct_of_peak_sustained_active': 'occ_pct',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed': 'sm_pct',
    'launch__registers_per_thread': 'regs',
    'launch__grid_size': 'grid',
    'launch__block_size': 'block',
    'smsp__inst_executed.sum': 'inst',
    'sm__cycles_active.avg': 'sm_cycles',
}
UNIT = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'Tbyte': 1e12,
        'ns': 1e-9, 'us': 1e-6, 'ms': 1e-3, 's': 1, 'nsecond': 1e-9, 'usecond': 1e-6, 'msecond': 1e-3, 'second': 1}


def rows_of(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = {}
        for h, u, v in zip(hdr, units, r):
            if h == 'Kernel Name':
                d['name'] = v.split('(')[0][-48:]
            elif h in KEYS:
                try:
                    d[KEYS[h]] = float(v.replace(',', '')) * UNIT.get(u, 1)
                except ValueError:
                    pass
            elif '--stalls' in sys.argv and h.startswith('smsp__average_warps_issue_stalled') and h.endswith('_per_issue_active.ratio'):
                try:
                    d.setdefault('stalls', {})[h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]] = float(v)
                except ValueError:
                    pass
        yield d


for path in [a for a in sys.argv[1:] if not a.startswith('--')]:
    print('==', path)
    for d in rows_of(path):
        tr = d.get('dram_rd', 0) + d.get('dram_wr', 0)
        print(f"{d.get('name', '?'):48s} {d.get('dur', 0) * 1e6:9.1f} us  dram {tr / 1e6:9.1f} MB ({d.get('dram_pct', 0):5.1f}%)"
              f"  tensor {d.get('tensor_pct', 0):5.1f}%  occ {d.get('occ_pct', 0):5.1f}%  sm {d.get('sm_pct', 0):5.1f}%"
              f"  regs {int(d.get('regs', 0))} grid {int(d.get('grid', 0))}x{int(d.get('block', 0))}")
        if 'stalls' in d:
            top = sorted(d['stalls'].items(), key=lambda kv: -kv[1])[:6]
            print('      stalls/issue: ' + ', '.join(f'{k} {v:.2f}' for k, v in top))
