#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel:
launch count, total and mean duration, share of the listed GPU time.
Usage: tools/launch_summary.py launches.csv [skip_first_n_launches]"""
import csv
import re
import sys
from collections import OrderedDict


def main():
    path = sys.argv[1]
    skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rows = []
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        name = re.sub(r"^void ", "", name)
        rows.append((int(r["ID"]), name, float(r["Metric Value"].replace(",", "")), r["Grid Size"], r["Block Size"]))
    rows = [r for r in rows if r[0] >= skip]
    agg = OrderedDict()
    for _, name, ns, grid, block in rows:
        a = agg.setdefault(name, [0, 0.0, grid, block])
        a[0] += 1
        a[1] += ns
    total = sum(a[1] for a in agg.values())
    print(f"# {path}: {len(rows)} launches (IDs >= {skip}), {total / 1e6:.3f} ms of GPU time listed")
    print(f"{'kernel':60s} {'launches':>8s} {'total ms':>11s} {'mean us':>11s} {'share':>7s}  grid / block (first launch)")
    for name, (n, ns, grid, block) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{name[:60]:60s} {n:8d} {ns / 1e6:11.3f} {ns / n / 1e3:11.1f} {100 * ns / total:6.1f}%  {grid} / {block}")


if __name__ == "__main__":
    main()
