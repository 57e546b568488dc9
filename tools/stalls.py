#!/usr/bin/env python3
"""stall breakdown per kernel from `ncu --page raw --csv`, plus top stall sites with context
from `--page source --csv --print-source sass`.  usage: stalls.py raw.csv sass.csv [filter] [nsites] [ctx]"""
import csv, sys
sys.path.insert(0, __file__.rsplit('/', 1)[0])
from sass_hot import kernels
raw, sass = sys.argv[1], sys.argv[2]
filt = sys.argv[3] if len(sys.argv) > 3 else ''
ns = int(sys.argv[4]) if len(sys.argv) > 4 else 6
ctx = int(sys.argv[5]) if len(sys.argv) > 5 else 5
rows = list(csv.reader(open(raw)))
hdr = rows[0]
names = [h for h in hdr if 'issue_stalled' in h and 'ratio' in h and 'not_issued' not in h]
for r in rows[2:]:
    if filt not in r[hdr.index('Kernel Name')]: continue
    print(r[hdr.index('Kernel Name')][:60], 'us', r[hdr.index('gpu__time_duration.sum')],
          'regs', r[hdr.index('launch__registers_per_thread')])
    vals = []
    for h in names:
        try: vals.append((float(r[hdr.index(h)].replace(',', '')), h))
        except Exception: pass
    print('   ' + '  '.join('%s %.2f' % (h.split('stalled_')[1].split('_per_')[0], v) for v, h in sorted(vals, reverse=True)[:8]))
for name, hdr, rows in kernels(sass):
    if filt not in name: continue
    ix = {h: i for i, h in enumerate(hdr)}
    S = lambda r: int(float(r[ix['# Samples']] or 0))
    E = lambda r: int(float(r[ix['Instructions Executed']] or 0))
    print(name[:70], 'samples', sum(S(r) for r in rows), 'exec', sum(E(r) for r in rows))
    order = sorted(range(len(rows)), key=lambda i: -S(rows[i]))[:ns]
    for i in sorted(order):
        print('-----')
        for j in range(max(0, i - ctx), i + 2):
            r = rows[j]; print(f"{E(r):7d} {S(r):5d}  {r[ix['Source']][:100]}")
