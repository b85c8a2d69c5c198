"""Summarise an ncu launch list of scripts/rn_steps.py (metrics gpu__time_duration.sum,
dram__bytes_read.sum, dram__bytes_write.sum): the last step, one line per launch, with the DRAM
bandwidth each launch reached; then totals per kernel name."""
import csv
import re
import sys
from collections import defaultdict


def main(path, show=True):
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    rows = list(csv.DictReader(lines))
    launches = {}
    order = []
    for x in rows:
        i = int(x["ID"])
        if i not in launches:
            launches[i] = {"name": x["Kernel Name"], "grid": x["Grid Size"]}
            order.append(i)
        v = float(x["Metric Value"].replace(",", ""))
        unit = x["Metric Unit"]
        if x["Metric Name"] == "gpu__time_duration.sum":
            launches[i]["us"] = v / 1000 if unit in ("ns", "nsecond") else v * {"us": 1, "usecond": 1, "ms": 1000, "msecond": 1000}.get(unit, 1)
        else:
            scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
            launches[i]["read" if "read" in x["Metric Name"] else "write"] = v * scale
    seq = [launches[i] for i in order]
    starts = [k for k, l in enumerate(seq) if "to_bf16_kernel" in l["name"]]
    last = seq[starts[-1]:]

    def short(n):
        n = n.replace("tfg::", "").replace("rn::", "").replace("tc::", "").replace("void ", "")
        return n.split("(")[0]

    agg = defaultdict(lambda: [0, 0.0, 0.0])
    for k, l in enumerate(last):
        b = l.get("read", 0) + l.get("write", 0)
        nm = short(l["name"])
        agg[nm][0] += 1
        agg[nm][1] += l["us"]
        agg[nm][2] += b
        if show:
            print("%4d %-44s %-14s %8.1f us %8.1f MB %6.2f TB/s" % (k, nm[:44], l["grid"], l["us"], b / 1e6, b / l["us"] / 1e6))
    print()
    tot = sum(v[1] for v in agg.values())
    for nm, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-48s %4d %8.3f ms %5.1f%% %9.1f MB %6.2f TB/s" % (nm[:48], v[0], v[1] / 1e3, 100 * v[1] / tot, v[2] / 1e6, v[2] / v[1] / 1e6))
    print("total %.3f ms" % (tot / 1e3))


if __name__ == "__main__":
    main(sys.argv[1], show=len(sys.argv) < 3)
