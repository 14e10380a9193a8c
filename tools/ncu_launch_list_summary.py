#!/usr/bin/env python
"""Per-kernel totals and shares of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv`).
    python tools/ncu_launch_list_summary.py launches.csv > summary.csv"""
import collections
import csv
import re
import sys


def main():
    rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
    h = rows[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        m = re.search(r"(?:<unnamed>::|at::native::|at::)?(\w+)(?:<[^(]*>)?\(", r[ki].replace("void ", ""))
        name = m.group(1) if m else r[ki][:40]
        v = float(r[vi].replace(",", "")) * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(r[ui], 1)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print("kernel,launches,total_us,avg_us,share_pct")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k},{n},{t:.1f},{t / n:.2f},{100 * t / tot:.1f}")


if __name__ == "__main__":
    main()
