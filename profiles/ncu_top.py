#!/usr/bin/env python
"""Top SASS instructions by warp-stall samples from `ncu --page source --csv` output (stdin)."""
import csv, sys
rows = list(csv.reader(sys.stdin))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
h = rows[hi]
end = next((i for i in range(hi + 1, len(rows)) if rows[i] and rows[i][0] in ('Address', 'Kernel Name')), len(rows))
body = [r for r in rows[hi + 1:end] if len(r) == len(h)]
si, so = h.index('# Samples'), h.index('Source')
stall_cols = [i for i, n in enumerate(h) if n.startswith('stall_') and 'Not Issued' not in n]
tot = sum(int(r[si] or 0) for r in body)
print('kernel:', rows[0][1][:120] if rows[0][0] == 'Kernel Name' else '?', 'total samples', tot)
agg = {}
for r in body:
    for i in stall_cols:
        agg[h[i]] = agg.get(h[i], 0) + int(r[i] or 0)
print('stall totals:', sorted(((v, k) for k, v in agg.items() if v), reverse=True)[:8])
n = int(sys.argv[1]) if len(sys.argv) > 1 else 25
for r in sorted(body, key=lambda r: -int(r[si] or 0))[:n]:
    st = sorted(((int(r[i] or 0), h[i][6:]) for i in stall_cols), reverse=True)[:2]
    print(f'{int(r[si]):7d} {100*int(r[si])/max(tot,1):5.1f}%  {r[so][:90]:90s} {st}')
