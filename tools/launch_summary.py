"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: python tools/launch_summary.py file.csv [n_detail]"""
import collections
import csv
import re
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
det = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
n = 0
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(row["Metric Value"].replace(",", ""))
    v = v / 1000 if row["Metric Unit"] == "ns" else (v * 1000 if row["Metric Unit"] == "ms" else v)
    k = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").replace("m2f::", "")
    agg[k][0] += 1
    agg[k][1] += v
    det[(k, row.get("Grid Size", ""))][0] += 1
    det[(k, row.get("Grid Size", ""))][1] += v
    tot += v
    n += 1
print(f"{n} launches, {tot:.1f} us total (serialised, cold cache)")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{t:9.1f} us {c:4d} x {t / c:7.1f} us {100 * t / tot:5.1f}%  {k[:90]}")
nd = int(sys.argv[2]) if len(sys.argv) > 2 else 0
for (k, g), (c, t) in sorted(det.items(), key=lambda kv: -kv[1][1])[:nd]:
    print(f"    {t:9.1f} us {c:4d} x {t / c:7.1f} us  {k[:60]} grid {g}")
