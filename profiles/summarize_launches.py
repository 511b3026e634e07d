#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name.
usage: python profiles/summarize_launches.py gpurun_out/launches.csv > profiles/launches_rNN.md"""
import csv
import re
import sys
from collections import OrderedDict


def main(path):
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            name = re.sub(r"\(.*", "", r["Kernel Name"])
            name = re.sub(r"^void ", "", name)
            rows.append((name, float(r["Metric Value"]), r["Grid Size"], r["Block Size"]))
    agg = OrderedDict()
    for n, ns, g, b in rows:
        a = agg.setdefault(n, [0, 0.0, g, b])
        a[0] += 1
        a[1] += ns
    tot = sum(a[1] for a in agg.values())
    print("| kernel | launches | total us | avg us | share | grid | block |")
    print("|---|---:|---:|---:|---:|---|---|")
    for n, (c, ns, g, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| `%s` | %d | %.1f | %.2f | %.1f%% | %s | %s |" % (n[:90], c, ns / 1e3, ns / 1e3 / c, 100 * ns / tot, g, b))
    print("\ntotal: %d launches, %.1f us (cold-cache, serialised under ncu: compare shares, not absolutes)" % (len(rows), tot / 1e3))


if __name__ == "__main__":
    main(sys.argv[1])
