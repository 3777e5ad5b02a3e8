"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv
import re
import sys
from collections import defaultdict

path = sys.argv[1]
rows = []
with open(path, newline="") as fh:
    lines = [l for l in fh if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    if unit == "us":
        v *= 1e3
    elif unit == "ms":
        v *= 1e6
    rows.append((r["Kernel Name"], v, r.get("Grid Size", ""), r.get("Block Size", "")))
agg = defaultdict(lambda: [0, 0.0])
for name, v, *_ in rows:
    short = re.sub(r"\(.*", "", name)
    agg[short][0] += 1
    agg[short][1] += v
total = sum(v for _, v in agg.values())
print(f"total {total/1e6:.3f} ms over {len(rows)} launches")
for name, (cnt, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v/1e6:9.3f} ms  {100*v/total:5.1f}%  n={cnt:5d}  avg={v/cnt/1e3:8.1f} us  {name}")
if len(sys.argv) > 2:
    print("\nslowest launches:")
    for name, v, g, b in sorted(rows, key=lambda r: -r[1])[: int(sys.argv[2])]:
        print(f"{v/1e3:9.1f} us grid={g} block={b} {name[:90]}")
