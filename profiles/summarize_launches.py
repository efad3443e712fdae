"""Turns an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel summary
(launches, total, mean, share). Usage: python profiles/summarize_launches.py launches.csv > out.csv"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
h = rows[hdr]
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "")
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += float(r[vi].replace(",", "")) / 1000.0
tot = sum(t for _, t in agg.values())
print("kernel,launches,total_us,mean_us,share_pct")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"\"{k}\",{c},{t:.1f},{t / c:.2f},{100 * t / tot:.2f}")
