#!/usr/bin/env python
"""per-kernel mean/max duration of an `ncu --metrics gpu__time_duration.sum --csv` launch list"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[hi]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
d = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= vi:
        continue
    n = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("<unnamed>::", "")
    v = float(r[vi].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[ui], 1.0)
    d.setdefault(n, []).append(v)
tot = 0.0
for n, v in d.items():
    if "native" in n or "at_cuda" in n:
        continue
    print("%-50s n=%4d mean_us=%9.1f max=%9.1f" % (n[:50], len(v), sum(v) / len(v), max(v)))
