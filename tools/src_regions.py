#!/usr/bin/env python3
"""From `ncu --page source --csv --print-source sass`: for kernels matching a filter, the hottest SASS windows
(samples per window of W lines) and the top lines with every non-zero stall-reason column.
usage: src_regions.py sass.csv [filter] [window] [ntop]"""
import sys
sys.path.insert(0, __file__.rsplit('/', 1)[0])
from sass_hot import kernels

path = sys.argv[1]
filt = sys.argv[2] if len(sys.argv) > 2 else ''
W = int(sys.argv[3]) if len(sys.argv) > 3 else 100
ntop = int(sys.argv[4]) if len(sys.argv) > 4 else 40
for name, hdr, rows in kernels(path):
    if filt not in name:
        continue
    ix = {h: i for i, h in enumerate(hdr)}
    def f(r, h):
        try:
            return float(r[ix[h]] or 0)
        except Exception:
            return 0.0
    S = [f(r, '# Samples') for r in rows]
    E = [f(r, 'Instructions Executed') for r in rows]
    tot = sum(S) or 1
    print('=' * 100)
    print(name[:90], 'lines', len(rows), 'samples', int(tot), 'exec', int(sum(E)))
    print('columns:', [h for h in hdr if 'stall' in h.lower()][:40])
    print('-- windows of %d lines: first line, samples %%, exec' % W)
    wins = []
    for i in range(0, len(rows), W):
        wins.append((sum(S[i:i + W]), i, sum(E[i:i + W])))
    for s, i, e in sorted(wins, reverse=True)[:25]:
        print('  line %6d  %5.2f%%  exec %11d  ipc-ish %.3f  %s' % (i, 100 * s / tot, int(e), e / max(s, 1), rows[i][ix['Source']][:60]))
    stall_cols = [h for h in hdr if h.lower().startswith('stall') or 'stall_' in h.lower()]
    print('-- top lines')
    for i in sorted(range(len(rows)), key=lambda i: -S[i])[:ntop]:
        r = rows[i]
        reasons = sorted(((f(r, h), h) for h in stall_cols if f(r, h) > 0), reverse=True)[:4]
        print('  line %6d %6d %5.2f%% exec %9d  %-60s %s' % (i, S[i], 100 * S[i] / tot, int(E[i]), r[ix['Source']][:60],
              ' '.join('%s=%d' % (h.replace('stall_', ''), v) for v, h in reasons)))
