"""Aggregates an ncu `--metrics gpu__time_duration.sum --csv` launch list per kernel name.
usage: python tools/launch_summary.py launches.csv [top]"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
gi = hdr.index("Grid Size") if "Grid Size" in hdr else None
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[ki])
    name = re.sub(r"mgb::|void ", "", name)[:100]
    v = float(r[vi].replace(",", ""))
    v = {"ns": v / 1e3, "us": v, "ms": v * 1e3, "s": v * 1e6}.get(r[ui], v)
    big = ""
    if gi is not None:
        g = eval(r[gi].replace("(", "").replace(")", "").split(",")[0])
        big = " [grid>=1000]" if g >= 1000 else ""
    agg[name + big][0] += 1
    agg[name + big][1] += v
tot = sum(v[1] for v in agg.values())
print("total %.1f us in %d launches" % (tot, sum(v[0] for v in agg.values())))
for k, v in sorted(agg.items(), key=lambda x: -x[1][1])[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    print("%6d %10.1f us %5.1f%%  %s" % (v[0], v[1], 100 * v[1] / tot, k))
