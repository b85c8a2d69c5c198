"""Summarise an ncu `--csv` launch list (one or several `--metrics`): per kernel count, mean duration,
share of the total, and -- when captured -- DRAM bytes and tensor-pipe activity per launch.
usage: python scripts/ncu_summary.py launches.csv [min_count]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
min_count = int(sys.argv[2]) if len(sys.argv) > 2 else 1
hdr = rows[0]
ki, mi, vi, ui, idi = (hdr.index(k) for k in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit", "ID"))
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
per, names = collections.defaultdict(dict), {}
for r in rows[1:]:
    v, u, m = float(r[vi].replace(",", "")), r[ui], r[mi]
    if m == "gpu__time_duration.sum":
        v = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
    elif m.startswith("dram__bytes"):
        v *= SCALE.get(u, 1)
    per[r[idi]][m] = v
    names[r[idi]] = r[ki].split("(")[0].replace("void ", "").replace("tfg::", "")[:64]
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0, 0.0])
for i, m in per.items():
    a = agg[names[i]]
    a[0] += 1
    a[1] += m.get("gpu__time_duration.sum", 0.0)
    a[2] += m.get("dram__bytes_read.sum", 0.0)
    a[3] += m.get("dram__bytes_write.sum", 0.0)
    a[4] += m.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 0.0)
tot = sum(a[1] for a in agg.values() if a[0] >= min_count)
print("%-66s %5s %9s %6s %9s %9s %7s" % ("kernel", "count", "mean us", "share", "rd MB", "wr MB", "tensor%"))
for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    if a[0] < min_count:
        continue
    print("%-66s %5d %9.2f %5.1f%% %9.2f %9.2f %7.1f" % (n, a[0], a[1] / a[0], 100 * a[1] / tot, a[2] / a[0] / 1e6,
                                                        a[3] / a[0] / 1e6, a[4] / a[0]))
print("total %.1f us over %d launches" % (tot, sum(a[0] for a in agg.values() if a[0] >= min_count)))
