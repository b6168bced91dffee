"""Top warp-stall instructions of a kernel in an ncu report (needs -lineinfo + --import-source on):
    python tools/stall_top.py <report.ncu-rep> <kernel regex> [n]
Prints the stall-reason mix, then the n SASS instructions with the most samples, with the CUDA source line when the report has it."""
import collections
import csv
import re
import subprocess
import sys

REASONS = ['stall_barrier', 'stall_branch_resolving', 'stall_dispatch', 'stall_lg', 'stall_long_sb', 'stall_math', 'stall_mio', 'stall_no_inst',
           'stall_not_selected', 'stall_selected', 'stall_short_sb', 'stall_wait']


def load(rep, kernel, view):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', view, '--kernel-name', 'regex:' + kernel,
                          '--launch-count', '1'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    data = []
    for r in rows[2:]:
        if len(r) < len(hdr) or r[0] == hdr[0] or r[0].startswith('Kernel Name'):
            break
        data.append(r)
    return hdr, data


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    hdr, data = load(rep, kernel, 'sass')
    idx = {h: i for i, h in enumerate(hdr)}
    S, SRC = idx['# Samples'], idx['Source']
    tot = sum(int(r[S]) for r in data)
    mix = collections.Counter()
    for r in data:
        for k in REASONS:
            mix[k] += int(r[idx[k]] or 0)
    s = sum(mix.values())
    print('%d SASS instructions, %d samples; stall mix: %s' % (len(data), tot, {k: round(100 * v / s, 1) for k, v in mix.most_common(8)}))
    for i in sorted(range(len(data)), key=lambda i: -int(data[i][S]))[:n]:
        r = data[i]
        c = {k: int(r[idx[k]] or 0) for k in REASONS}
        print('%5d %6s %4.1f%% %-18s %s' % (i, r[S], 100 * int(r[S]) / tot, max(c, key=c.get), r[SRC].strip()[:90]))
    try:
        hdr2, lines = load(rep, kernel, 'cuda')
        i2 = {h: i for i, h in enumerate(hdr2)}
        if '# Samples' in i2:
            print('-- by CUDA source line')
            for r in sorted(lines, key=lambda r: -int(r[i2['# Samples']] or 0))[:n]:
                print('%6s %4.1f%%  %s' % (r[i2['# Samples']], 100 * int(r[i2['# Samples']] or 0) / tot, r[i2['Source']].strip()[:120]))
    except Exception as e:      # reports without source
        print('(no CUDA source view: %s)' % e)


if __name__ == '__main__':
    main()
