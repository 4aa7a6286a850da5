"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv):  python tools/launch_summary.py FILE.csv"""
import collections
import csv
import re
import sys

rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]
ki, vi, ui, gi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit"), hdr.index("Grid Size")
mi = hdr.index("Metric Name")
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
for r in rows[1:]:
    if len(r) <= vi or r[vi] in ("n/a", "") or r[mi] != "gpu__time_duration.sum":
        continue
    v = float(r[vi].replace(",", ""))
    v = v / 1000 if r[ui] == "ns" else v * 1000 if r[ui] == "ms" else v
    key = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("unnamed>::", "") + " grid " + r[gi]
    agg[key][0] += 1
    agg[key][1] += v
    tot += v
print("%d launches, %.1f us in total" % (sum(c for c, _ in agg.values()), tot))
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-72s %5d x %8.2f us = %9.1f us  %5.1f %%" % (n[:72], c, t / c, t, 100 * t / tot))
