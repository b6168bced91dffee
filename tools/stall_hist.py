"""Stall-sample histogram of one kernel from `ncu -i rep --page source --csv` output (SASS view).
    ncu -i prof.ncu-rep --page source --csv --kernel-name regex:NAME --launch-count 1 > src.csv;  python tools/stall_hist.py src.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h0 = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr, data = rows[h0], [r for r in rows[h0 + 1:] if len(r) > 5 and r[0] != 'Address']
ix = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ix['# Samples']]) for r in data)
print('total samples', tot, 'instructions', len(data))
stall_cols = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.006
for k, r in enumerate(data):
    n = int(r[ix['# Samples']])
    if n > tot * thr:
        top = sorted(((int(r[ix[c]]), c) for c in stall_cols), reverse=True)[:2]
        print(k, n, '%.1f%%' % (100 * n / tot), r[ix['Source']].strip()[:70], top)
byop = {}
for r in data:
    t = r[ix['Source']].split()
    op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    byop[op] = byop.get(op, 0) + int(r[ix['# Samples']])
print([(k, '%.1f%%' % (100 * v / tot)) for k, v in sorted(byop.items(), key=lambda x: -x[1])[:20]])
print({c: '%.1f%%' % (100 * sum(int(r[ix[c]]) for r in data) / tot) for c in stall_cols})
