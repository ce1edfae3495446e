#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: total time and share per kernel name.

    python scripts/summarize_launches.py gpurun_out/launches.csv [header lines ...] > profiles/rN_launches_summary.txt
"""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    rows = [l for l in open(path) if l.startswith('"')]
    rd = csv.DictReader(rows)
    tot = collections.defaultdict(float)
    cnt = collections.Counter()
    grids = collections.defaultdict(set)
    for r in rd:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        v = float(r["Metric Value"].replace(",", ""))
        if r["Metric Unit"] in ("us", "usecond"):
            v *= 1e3
        tot[name] += v
        cnt[name] += 1
        grids[name].add((r["Grid Size"], r["Block Size"]))
    total = sum(tot.values())
    for h in sys.argv[2:]:
        print("# " + h)
    print(f"total {total / 1e6:.3f} ms over {sum(cnt.values())} launches")
    for name, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        g = sorted(grids[name])[0]
        print(f"{v / 1e6:9.3f} ms {100 * v / total:5.1f}% n={cnt[name]:4d} avg={v / cnt[name] / 1e3:9.1f} us  {name[:70]}  grid{g[0]} block{g[1]}")


if __name__ == "__main__":
    main()
