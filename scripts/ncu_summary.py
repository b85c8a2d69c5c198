"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel count,
mean duration and share of the total.  usage: python scripts/ncu_summary.py launches.csv"""
import collections
import csv
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] == "ns" else v * 1e3 if r[ui] == "ms" else v
    name = r[ki].split("(")[0].replace("void ", "").replace("tfg::", "")[:70]
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print("%-72s %5s %10s %7s" % ("kernel", "count", "mean us", "share"))
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-72s %5d %10.2f %6.1f%%" % (n, c, t / c, 100 * t / tot))
print("total %.1f us over %d launches" % (tot, sum(v[0] for v in agg.values())))
