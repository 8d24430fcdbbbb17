"""Print selected metrics of every launch in an `ncu --page raw --csv` dump: python tools/ncu_raw.py dump.csv metric..."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = rows[0]
keys = sys.argv[2:]
for i, c in enumerate(h):
    if c in keys or any(k.endswith("*") and c.startswith(k[:-1]) for k in keys):
        print(c, rows[1][i], [r[i] for r in rows[2:]])
