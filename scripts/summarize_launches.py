#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel totals.

usage: summarize_launches.py launches.csv [title] > profiles/launch_breakdown_xxx.txt
"""
import csv
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else path
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
        rows.append((r["Kernel Name"], us))
    tot = sum(u for _, u in rows)
    agg = defaultdict(lambda: [0.0, 0])
    for k, u in rows:
        agg[k][0] += u
        agg[k][1] += 1
    print(title)
    print(f"total us {tot:.0f}  launches {len(rows)}")
    for k, (u, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{u:10.0f} us {100 * u / tot:5.1f}%  n={n:4d}  avg {u / n:8.1f} us  {k[:70]}")


if __name__ == "__main__":
    main()
