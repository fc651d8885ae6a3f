#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (the text files under
profiles/*_ncu_launches_summary.txt).   usage: python profiles/launch_summary.py launches.csv [skip_first_n]"""
import csv
import collections
import re
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
hdr, data = rows[0], rows[1:]
ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
data = data[skip:]
tot = collections.Counter()
cnt = collections.Counter()
for r in data:
    name = re.sub(r"^void |unnamed>::|mcb::|\(.*$", "", r[ik])
    tot[name] += float(r[iv].replace(",", ""))
    cnt[name] += 1
total = sum(tot.values())
print(f"total kernel time: {total / 1e6:.3f} ms over {len(data)} launches")
print(f"{'kernel':60s} {'launches':>8s} {'total ms':>10s} {'avg us':>10s} {'share':>7s}")
for k, v in tot.most_common():
    print(f"{k:60s} {cnt[k]:8d} {v / 1e6:10.3f} {v / cnt[k] / 1e3:10.1f} {100 * v / total:6.1f}%")
