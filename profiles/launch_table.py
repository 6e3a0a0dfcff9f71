#!/usr/bin/env python3
"""Aggregates an ncu launch list (`--metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --csv`)
by kernel, split into the prefill and ONE decode step of a generation run; with the DRAM metrics it also prints the
step's measured DRAM traffic and (--json FILE ARCH BATCH) writes it for bench.py's roofline.traffic.
usage: python profiles/launch_table.py launches.csv [--json out.json arch batch]"""
import collections
import csv
import json
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
iK, iV, iG, iM, iI = (hdr.index(c) for c in ('Kernel Name', 'Metric Value', 'Grid Size', 'Metric Name', 'ID'))
launch = collections.OrderedDict()                    # launch id -> [name, grid, {metric: value}]
for r in rows[1:]:
    e = launch.setdefault(r[iI], [r[iK], r[iG], {}])
    e[2][r[iM]] = float(r[iV].replace(',', ''))
data = list(launch.values())
TIME, RD, WR = 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum'
# step boundaries: the kernel that opens a step on the classic path, the one that closes it on the deferred-norm path
opens = [i for i, (n, _, _) in enumerate(data) if 'advance_step' in n]
closes = [i for i, (n, _, _) in enumerate(data) if 'decode_tail' in n]
if opens:
    pre_end, a, b = opens[0], opens[0], (opens[1] if len(opens) > 1 else len(data))
elif len(closes) >= 2:
    pre_end, a, b = closes[0] + 1, closes[0] + 1, closes[1] + 1
else:
    pre_end, a, b = len(data), len(data), len(data)


def short(n):
    m = re.search(r'(\w+)<(.*?)>\(', n)
    return (m.group(1) + '<' + m.group(2) + '>') if m else n.split('(')[0][-40:]


traffic = None
for label, (lo, hi) in {'prefill': (0, pre_end), 'decode step': (a, b)}.items():
    agg = collections.OrderedDict()
    for n, g, m in data[lo:hi]:
        e = agg.setdefault((short(n), g), [0, 0.0, 0.0])
        e[0] += 1
        e[1] += m.get(TIME, 0.0)
        e[2] += m.get(RD, 0.0) + m.get(WR, 0.0)
    tot = sum(v[1] for v in agg.values()) or 1.0
    byt = sum(v[2] for v in agg.values())
    print(f'== {label}: {hi - lo} launches, {tot / 1e3:.1f} us (cold-cache, serialised: compare shares)' +
          (f', DRAM traffic {byt / 1e6:.1f} MB' if byt else ''))
    for k, (c, v, by) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
        print(f'  {k[0]:48s} grid {k[1]:16s} x{c:3d} {v / 1e3:9.1f} us {100 * v / tot:5.1f}%  avg {v / c / 1e3:8.2f} us' +
              (f'  {by / c / 1e6:8.2f} MB per launch' if byt else ''))
    if label == 'decode step':
        traffic = byt
if '--json' in sys.argv and traffic:
    i = sys.argv.index('--json')
    json.dump({'arch': sys.argv[i + 2], 'batch': int(sys.argv[i + 3]), 'dram_bytes_per_step': traffic,
               'how': 'ncu dram__bytes_read.sum + dram__bytes_write.sum summed over the launches of one decode step '
                      '(profiles/launch_table.py)'}, open(sys.argv[i + 1], 'w'))
