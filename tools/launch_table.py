"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: avg device time per kernel."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
d = collections.OrderedDict()
for r in rows[hdr + 1:]:
    rec = dict(zip(rows[hdr], r))
    if rec.get("Metric Name") == "gpu__time_duration.sum":
        d.setdefault(rec["Kernel Name"][:70], []).append(float(rec["Metric Value"].replace(",", "")))
tot = 0
for k, v in d.items():
    a = sum(v) / len(v) / 1000
    tot += a
    print(f"{k:72s} n={len(v):3d} avg={a:6.2f} us")
print(f"{'sum':72s}       {tot:6.2f} us")
