"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel name count, total and share."""
import csv
import re
import sys
from collections import defaultdict

rows = []
with open(sys.argv[1], newline="") as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}.get(unit, 1e-3)
    name = r["Kernel Name"]
    name = re.sub(r"\(.*$", "", name)
    rows.append((name, v * scale))
agg = defaultdict(lambda: [0, 0.0])
for n, t in rows:
    agg[n][0] += 1
    agg[n][1] += t
tot = sum(t for _, t in rows)
print("launches %d   total %.1f us" % (len(rows), tot))
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%6.2f %%  %9.1f us  %6d x %8.2f us   %s" % (100 * t / tot, t, c, t / c, n[:110]))
