#!/usr/bin/env python
"""ncu launch list (--metrics gpu__time_duration.sum[,sm__cycles_elapsed.max] --csv) -> per-kernel summary CSV:
launches, total us, mean us, share of all kernel time, mean SM cycles.   usage: launch_summary.py in.csv out.csv"""
import csv
import re
import sys
from collections import OrderedDict


def main(src, dst):
    lines = [l for l in open(src) if l.startswith('"')]
    rows = list(csv.DictReader(lines))
    agg = OrderedDict()
    for r in rows:
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
        a = agg.setdefault(name, {"n": 0, "ns": 0.0, "cyc": 0.0, "grid": r["Grid Size"], "block": r["Block Size"]})
        v = float(r["Metric Value"].replace(",", ""))
        if r["Metric Name"] == "gpu__time_duration.sum":
            a["n"] += 1; a["ns"] += v
        elif r["Metric Name"] == "sm__cycles_elapsed.max":
            a["cyc"] += v
    tot = sum(a["ns"] for a in agg.values())
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel", "grid", "block", "launches", "total_us", "mean_us", "share_of_kernel_time", "mean_sm_cycles"])
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ns"]):
            w.writerow([k, a["grid"], a["block"], a["n"], round(a["ns"] / 1e3, 1), round(a["ns"] / 1e3 / max(a["n"], 1), 2),
                        round(a["ns"] / tot, 4), round(a["cyc"] / max(a["n"], 1))])


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
