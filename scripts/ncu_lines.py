"""Per-CUDA-source-line dynamic instruction counts and stall samples of an .ncu-rep (read here, no GPU).
usage: ncu_lines.py report.ncu-rep [units] [top]   (units: divide counts by this, e.g. systems per launch)"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
N = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 35
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
agg, fname = [], None
for r in csv.reader(io.StringIO(txt)):
    if len(r) == 2 and r[0] in ("File Path", "File Name"):
        fname = r[1].split("/")[-1]
        continue
    if len(r) < 10 or r[0] == "Line No" or r[0] == "":
        continue
    try:
        agg.append((fname, int(r[0]), r[1].strip(), int(r[7]), int(r[4] or 0)))
    except ValueError:
        pass
tot = sum(a[3] for a in agg)
ts = sum(a[4] for a in agg) or 1
print("warp instructions per unit: %.1f" % (tot / N))
agg.sort(key=lambda x: -x[3])
for a in agg[:top]:
    print("%-16s %4d %9.1f %5.1f%%smp  %s" % (a[0], a[1], a[3] / N, 100.0 * a[4] / ts, a[2][:95]))
