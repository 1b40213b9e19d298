#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections
import csv
import re
import sys


def main(path, title=""):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg, tot, n = {}, 0.0, 0
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1000 if u in ("nsecond", "ns") else (v * 1000 if u in ("msecond", "ms") else v)
        name = re.sub(r"\(.*", "", row["Kernel Name"])[:80]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        tot += v
        n += 1
    print(f"# {title}\n\n{n} launches, {tot:.0f} us total (serialised launches; cold L2 unless the title says --cache-control none: compare shares)\n")
    print("| kernel | launches | total us | us/launch | share |\n|---|---:|---:|---:|---:|")
    for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"| `{k}` | {c} | {t:.1f} | {t / c:.1f} | {100 * t / tot:.1f}% |")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
