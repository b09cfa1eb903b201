#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name (times are cold-cache and
serialised under ncu: compare SHARES, not absolutes).  usage: summarize_launches.py launches.csv [top]"""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot, n = 0.0, 0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"^void ", "", re.sub(r"\(.*", "", row["Kernel Name"]))
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e6 if u == "ns" else v / 1e3 if u == "us" else v
        agg[name][0] += 1
        agg[name][1] += v
        tot += v
        n += 1
    print(f"# {path}: {n} launches, {tot:.2f} ms total device time under ncu")
    print(f"# {'ms':>10} {'share':>7} {'launches':>8}  kernel")
    for k, (c, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"{ms:12.3f} {100 * ms / tot:6.2f}% {c:8d}  {k[:120]}")


if __name__ == "__main__":
    main()
