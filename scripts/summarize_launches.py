#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches,
total time and share, for the LAST `--steps` worth of launches (or everything).
usage: summarize_launches.py launches.csv [--tail-frac F]"""
import collections
import csv
import re
import sys


def short(name):
    name = re.sub(r"\(.*", "", name)
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"<.*", "", name)
    return name[-70:]


def main():
    path = sys.argv[1]
    frac = float(sys.argv[sys.argv.index("--tail-frac") + 1]) if "--tail-frac" in sys.argv else 1.0
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr, rows = rows[0], rows[1:]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    rows = rows[int(len(rows) * (1 - frac)):]
    tot = collections.defaultdict(float)
    cnt = collections.Counter()
    for r in rows:
        k = short(r[ki])
        tot[k] += float(r[vi].replace(",", "")) / 1e6
        cnt[k] += 1
    total = sum(tot.values())
    print(f"# {path}: {len(rows)} launches, {total:.2f} ms total (cold-cache, serialised under ncu)")
    print(f"{'kernel':72s} {'launches':>8s} {'ms':>10s} {'share':>7s}")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        print(f"{k:72s} {cnt[k]:8d} {v:10.3f} {v / total:7.3f}")


if __name__ == "__main__":
    main()
