#!/usr/bin/env python
"""Per-kernel durations from an `ncu --metrics gpu__time_duration.sum --csv` log: name, launches, last few times (us)."""
import collections
import csv
import sys

rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
h = rows[0]
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
d = collections.OrderedDict()
for r in rows[1:]:
    d.setdefault(r[ki][:70], []).append(float(r[vi].replace(",", "")) / 1000.0)
tail = int(sys.argv[2]) if len(sys.argv) > 2 else 4
for k, v in d.items():
    print(f"{k:72s} n={len(v):4d}  " + " ".join(f"{x:8.1f}" for x in v[-tail:]))
