"""Top stall-sample source lines / SASS from `ncu -i X.ncu-rep --page source --csv --print-source sass,cuda`.
Usage: python profiles/hot_lines.py src.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
h = rows[hi]
si = h.index("Warp Stall Sampling (All Samples)")
ii = h.index("Instructions Executed")
body = [r for r in rows[hi + 1:] if len(r) > si and r[si].replace(".", "").isdigit()]
tot = sum(float(r[si]) for r in body) or 1
by_line = {}
for r in body:
    k = (r[0], r[1].strip()[:110])
    by_line[k] = by_line.get(k, 0) + float(r[si])
print(f"total samples {tot:.0f}")
for (ln, src), v in sorted(by_line.items(), key=lambda kv: -kv[1])[:N]:
    print(f"{100 * v / tot:5.1f}%  L{ln}: {src}")
print("--- top SASS")
for r in sorted(body, key=lambda r: -float(r[si]))[:N]:
    print(f"{100 * float(r[si]) / tot:5.1f}%  L{r[0]} {r[3].strip()[:90]}  inst={r[ii]}")
