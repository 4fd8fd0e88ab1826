"""Summarises an `ncu --page source --csv` dump: stall reasons in total and the hottest SASS lines.
usage: python tools/ncu_stalls.py src.csv [top]"""
import csv
import sys
import collections

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
h = None
for i, r in enumerate(rows):
    if r and r[0] == "Address":
        h = i
        break
hdr = rows[h]
idx = {k: i for i, k in enumerate(hdr)}
stalls = [k for k in hdr if k.startswith("stall_") and "Not Issued" not in k]
tot = collections.Counter()
lines = []
ninst = 0
for r in rows[h + 1:]:
    if len(r) < len(hdr) or not r[0].startswith("0x"):
        continue
    s = int(r[idx["# Samples"]] or 0)
    ninst += int(r[idx["Instructions Executed"]] or 0)
    for k in stalls:
        tot[k] += int(r[idx[k]] or 0)
    lines.append((s, r[idx["Source"]].strip(), {k: int(r[idx[k]] or 0) for k in stalls if int(r[idx[k]] or 0) > 0},
                  int(r[idx["Instructions Executed"]] or 0)))
all_s = sum(tot.values())
print("warp instructions executed: %d ; samples %d" % (ninst, all_s))
for k, v in tot.most_common(10):
    print("  %-28s %6.1f%%" % (k, 100.0 * v / max(all_s, 1)))
ops = collections.Counter()
for s, src, d, n in lines:
    ops[src.split()[0] if not src.startswith("@") else src.split()[1]] += n
print("instruction mix:", ", ".join("%s %.1f%%" % (k, 100.0 * v / ninst) for k, v in ops.most_common(14)))
print("hottest lines:")
for i, (s, src, d, n) in enumerate(sorted(lines, key=lambda x: -x[0])[:top]):
    print("  %5d  %-60s %s" % (s, src[:60], sorted(d.items(), key=lambda x: -x[1])[:3]))
