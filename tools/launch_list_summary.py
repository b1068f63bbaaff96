#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel name, launches, total time
and share of all profiled kernel time.  usage: launch_list_summary.py launches.csv [title] > profiles/x.md"""
import csv
import re
import sys
from collections import OrderedDict


def main():
    path = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else path
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    rows = list(csv.DictReader(lines))
    agg = OrderedDict()
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"^void\s+", "", r["Kernel Name"])
        name = re.sub(r"\(.*\)$", "", name).replace("bvc::", "")
        ns = float(r["Metric Value"].replace(",", ""))
        if r["Metric Unit"] in ("us", "usecond"):
            ns *= 1e3
        elif r["Metric Unit"] in ("ms", "msecond"):
            ns *= 1e6
        a = agg.setdefault(name, [0, 0.0, r["Grid Size"], r["Block Size"]])
        a[0] += 1
        a[1] += ns
    total = sum(a[1] for a in agg.values())
    print("# %s\n" % title)
    print("%d kernel launches profiled, %.3f ms of kernel time in total (per-launch times under ncu are cold-cache and "
          "serialised: compare SHARES with bench.py's `kernels_ms`, not absolutes).\n" % (sum(a[0] for a in agg.values()), total / 1e6))
    print("| kernel | launches | total ms | mean us | share | grid (last) | block |")
    print("|---|---|---|---|---|---|---|")
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| %s | %d | %.3f | %.1f | %.1f %% | %s | %s |" % (name, a[0], a[1] / 1e6, a[1] / a[0] / 1e3, 100 * a[1] / total, a[2], a[3]))


if __name__ == "__main__":
    main()
