#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and mean time.
usage: launch_summary.py launches.csv [first_kernel_substr last_kernel_substr]"""
import collections
import csv
import re
import sys


def short(n):
    m = re.search(r"gemm_kernel<(\d+), (\d+), (\d+), (\d+), opb::(\w+)", n)
    if m:
        return f"gemm<BN={m.group(1)},st={m.group(2)},g={m.group(3)},split={m.group(4)},{m.group(5)}>"
    n = re.sub(r"^void ", "", n)
    n = re.sub(r"\(.*", "", n)
    n = re.sub(r"<.*", "", n)
    return n[:70]


def main():
    lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
    rows = list(csv.DictReader(lines))
    a, b = 0, len(rows)
    if len(sys.argv) >= 4:
        first = [i for i, d in enumerate(rows) if sys.argv[2] in d["Kernel Name"]]
        last = [i for i, d in enumerate(rows) if sys.argv[3] in d["Kernel Name"]]
        sel = int(sys.argv[4]) if len(sys.argv) > 4 else -1
        b = last[sel] + 1
        a = max(i for i in first if i < b) if sel == -1 else first[sel]
    agg = collections.OrderedDict()
    for d in rows[a:b]:
        n = short(d["Kernel Name"])
        agg.setdefault(n, [0, 0.0])
        agg[n][0] += 1
        agg[n][1] += float(d["Metric Value"])
    tot = sum(v[1] for v in agg.values())
    print(f"launches {b - a}, total {tot / 1e3:.1f} us")
    for n, (c, t) in agg.items():
        print(f"{t / 1e3:9.1f} us {100 * t / tot:5.1f}% {c:4d} x {t / c / 1e3:7.1f}  {n}")


if __name__ == "__main__":
    main()
