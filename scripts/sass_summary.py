#!/usr/bin/env python
"""Per-kernel SASS instruction counts of libb200mp.so (cuobjdump -sass): tensor-core MMAs (UTCHMMA), tensor-memory loads / stores
(LDTM / STTM), bulk copies (UBLKCP = cp.async.bulk, 1-D TMA), tensor-map TMA (UTMALDG / UTMASTG), vector REDs, barriers.
usage: python scripts/sass_summary.py > profiles/sass_summary.txt"""
import collections
import os
import re
import subprocess

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "models-for-relational-multimodal-data_b200", "libb200mp.so")
OPS = ["UTCHMMA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "REDG", "ATOMG", "BAR.SYNC", "SYNCS", "LDG", "STG", "LDS", "STS"]
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
counts, cur = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = demangle(m.group(1))
        cur = re.sub(r"\(.*", "", cur).replace("b200mp::", "")
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        counts[cur]["total"] += 1
        for o in OPS:
            if op == o or op.startswith(o + "."):
                counts[cur][o] += 1
print("# cuobjdump -sass libb200mp.so (sm_100a): instruction counts per kernel; UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st,")
print("# UBLKCP = cp.async.bulk (1-D TMA), UTMALDG/UTMASTG = tensor-map TMA (none: the edge tiles are row-gathered, see DESIGN.md 4.3)")
print(f"{'kernel':78s} {'total':>6s} " + " ".join(f"{o:>8s}" for o in OPS))
for k, c in counts.items():
    if c["total"]:
        print(f"{k[:78]:78s} {c['total']:6d} " + " ".join(f"{c[o]:8d}" for o in OPS))
