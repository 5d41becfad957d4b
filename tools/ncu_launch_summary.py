#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel time of ONE training step
(the launches between two consecutive AdamW groups).  Usage: ncu_launch_summary.py launches.csv [step_index]"""
import collections
import csv
import re
import sys


def load(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rows = []
    for x in csv.DictReader(lines):
        if x.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((re.sub(r"\(.*", "", x["Kernel Name"]), float(x["Metric Value"].replace(",", "")) / 1e3))
    return rows


def main():
    rows = load(sys.argv[1])
    which = int(sys.argv[2]) if len(sys.argv) > 2 else -1
    ends = [i for i, (n, _) in enumerate(rows) if "adamw" in n and (i + 1 == len(rows) or "adamw" not in rows[i + 1][0])]
    if len(ends) < 2:
        seg = rows
    else:
        a, b = (ends[which - 1] + 1, ends[which] + 1) if which != 0 else (0, ends[0] + 1)
        seg = rows[a:b]
    agg = collections.OrderedDict()
    for n, us in seg:
        c = agg.setdefault(n, [0, 0.0])
        c[0] += 1
        c[1] += us
    tot = sum(v[1] for v in agg.values())
    print(f"launches in step: {len(seg)}   sum of kernel durations: {tot / 1e3:.3f} ms (ncu-serialised, cold cache)")
    print(f"{'kernel':60s} {'n':>5s} {'ms':>9s} {'share':>7s} {'us/launch':>10s}")
    for n, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n[:60]:60s} {c:5d} {us / 1e3:9.3f} {us / tot:7.2%} {us / c:10.1f}")


if __name__ == "__main__":
    main()
