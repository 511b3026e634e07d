#!/usr/bin/env python
"""Top stall locations from `ncu -i X.ncu-rep --page source --csv --kernel-name regex:K > src.csv`."""
import csv
import sys


def main(path, n=40):
    rows = list(csv.reader(open(path)))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = rows[2:]

    def f(r, k):
        try:
            return float(r[ix[k]])
        except Exception:
            return 0.0
    tot = sum(f(r, "# Samples") for r in data) or 1.0
    print("kernel:", rows[0][1][:100])
    print("total samples %d, warp instructions executed %d" % (tot, sum(f(r, "Instructions Executed") for r in data)))
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = sorted(((sum(f(r, s) for r in data), s) for s in stalls), reverse=True)[:8]
    print("stall mix:", ", ".join("%s %.1f%%" % (s, 100 * v / tot) for v, s in agg))
    for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:n]:
        st = sorted([(f(r, s), s) for s in stalls], reverse=True)[:2]
        print("%5.1f%% ex=%8d %-72s %s" % (100 * f(r, "# Samples") / tot, f(r, "Instructions Executed"),
                                          r[ix["Source"]][:72], ", ".join("%s=%d" % (s, v) for v, s in st)))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
