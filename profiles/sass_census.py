#!/usr/bin/env python3
"""Per-kernel count of the SASS mnemonics that show which hardware path a kernel uses (run here, no GPU needed):
UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG/UTMAREDG = TMA tensor copies, UBLKCP = bulk copies
(incl. shared -> cluster shared), UTCBAR = tcgen05.commit, HMMA = mma.sync, LDGSTS = cp.async.
usage: python profiles/sass_census.py [libkx_b200.so] > profiles/r02_sass_census.txt"""
import collections
import os
import re
import subprocess
import sys

so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                         "ai-agent-kubectl_b200", "libkx_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
pat = re.compile(r"\b(UTC[A-Z]*MMA|LDTM|STTM|UTMALDG|UTMASTG|UTMAREDG|UTMAPF|UBLKCP|UBLKPF|UTCBAR|HMMA|LDGSTS|SYNCS|UCGABAR_ARV|REDG|ATOMG)\b")
counts, name = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(anonymous namespace\)::|kx::|<unnamed>::", "", name).split("(")[0]
        counts[name] = collections.Counter()
        continue
    if name:
        for k in pat.findall(line):
            counts[name][k] += 1
cols = ["UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "UTCBAR", "UCGABAR_ARV", "HMMA", "LDGSTS"]
print(f"{'kernel':<70} " + " ".join(f"{c:>11}" for c in cols))
tot = collections.Counter()
for n, c in counts.items():
    if not any(c[k] for k in cols):
        continue
    print(f"{n[:70]:<70} " + " ".join(f"{c[k]:>11}" for k in cols))
    tot.update(c)
print(f"{'TOTAL':<70} " + " ".join(f"{tot[k]:>11}" for k in cols))
