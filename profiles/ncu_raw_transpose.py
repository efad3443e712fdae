"""`ncu -i X.ncu-rep --page raw --csv` -> metric,unit,value rows of ONE launch (the profiles/*_ncu_full_*.csv format).
Usage: python profiles/ncu_raw_transpose.py X.ncu-rep [launch index] > out.csv"""
import csv
import subprocess
import sys

rep, idx = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2 + idx]
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit", "value"])
for h, u, v in zip(hdr, units, data):
    w.writerow([h, u, v])
