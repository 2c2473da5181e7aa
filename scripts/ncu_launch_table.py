"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per (kernel, grid): launches, total and mean device
time and the share of the listed time.   python scripts/ncu_launch_table.py launches.csv [top]"""
import collections
import csv
import re
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    k = re.sub(r"\(.*", "", row["Kernel Name"])
    v = float(row["Metric Value"].replace(",", ""))
    v = v / 1e3 if row["Metric Unit"] == "ns" else v * 1e3 if row["Metric Unit"] == "ms" else v
    a = agg.setdefault((k, row.get("Grid Size", "")), [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print("launches %d, listed device time %.1f us" % (sum(a[0] for a in agg.values()), tot))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
for (k, g), a in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
    print("%-44s %-14s n=%5d  sum=%10.1f us  mean=%8.2f us  %5.1f%%" % (k[:44], g, a[0], a[1], a[1] / a[0], 100 * a[1] / tot))
