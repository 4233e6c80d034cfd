#!/usr/bin/env python
"""Mean duration per kernel name from an ncu launch-list CSV (gpu__time_duration.sum). usage: launch_means.py <csv> [skip_first_n]"""
import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hdr]
kn, mv, mu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
acc = collections.defaultdict(list)
for r in rows[hdr + 1:]:
    if len(r) <= mv:
        continue
    v = float(r[mv].replace(",", ""))
    u = r[mu]
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
    acc[re.sub(r"\(.*", "", r[kn])].append(v)
for k, v in sorted(acc.items(), key=lambda kv: -sum(kv[1])):
    v = v[skip:] if len(v) > skip else v
    print(f"{k[:60]:60s} n {len(v):4d} mean {sum(v)/len(v):9.1f} us")
