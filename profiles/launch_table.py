#!/usr/bin/env python3
"""Aggregates an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel,
split into the prefill and the decode part of a generation run."""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
iK, iV, iG = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Grid Size')
data = [(r[iK], float(r[iV].replace(',', '')), r[iG]) for r in rows[1:]]
idx = [i for i, (n, _, _) in enumerate(data) if 'advance_step' in n] or [len(data)]


def short(n):
    m = re.search(r'(\w+)<(.*?)>\(', n)
    return (m.group(1) + '<' + m.group(2) + '>') if m else n.split('(')[0][-40:]


end = idx[1] if len(idx) > 1 else len(data)            # several steps captured: tabulate the first one
for label, (a, b) in {'prefill': (0, idx[0]), 'decode step': (idx[0], end)}.items():
    agg = collections.OrderedDict()
    for n, v, g in data[a:b]:
        k = (short(n), g)
        agg.setdefault(k, [0, 0.0])
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v for _, v in agg.values())
    print(f'== {label}: {b - a} launches, {tot / 1e3:.1f} us (cold-cache, serialised: compare shares)')
    for k, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
        print(f'  {k[0]:48s} grid {k[1]:16s} x{c:3d} {v / 1e3:9.1f} us {100 * v / tot:5.1f}%  avg {v / c / 1e3:8.2f} us')
