"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list: python tools/launch_table.py file.csv [last_n]"""
import collections
import csv
import sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
seq = []
for r in rows[1:]:
    v = float(r[vi].replace(',', ''))
    v = v / 1000 if r[ui] in ('ns', 'nsecond') else v * 1000 if r[ui] in ('ms', 'msecond') else v
    seq.append((r[ki][:70], v))
if len(sys.argv) > 2:
    seq = seq[-int(sys.argv[2]):]
agg = collections.OrderedDict()
for n, v in seq:
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1
    a[1] += v
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print('%-72s %5d %10.1f us %8.2f us each' % (n, c, t, t / c))
print('total %.1f us over %d launches' % (sum(v for _, v in seq), len(seq)))
