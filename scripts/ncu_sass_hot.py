#!/usr/bin/env python
"""Per-SASS-instruction executed counts + stall samples from an .ncu-rep source page.
usage: ncu_sass_hot.py rep [min_exec_fraction]   -> prints instructions whose exec count >= frac * max"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
lines = out.splitlines()
# first kernel only
start = [i for i, l in enumerate(lines) if l.startswith('"Address"')][0]
end = next((i for i in range(start + 1, len(lines)) if lines[i].startswith('"Kernel Name"')), len(lines))
rows = list(csv.reader(io.StringIO("\n".join(lines[start:end]))))
h = rows[0]
ie, ss, src = h.index("Instructions Executed"), h.index("# Samples"), h.index("Source")
data = [(r[src].strip(), int(r[ie]), int(r[ss])) for r in rows[1:] if len(r) > ie]
mx = max(d[1] for d in data)
tot = sum(d[1] for d in data)
print("total inst", tot, "max per-instr exec", mx, "n instr", len(data))
mode = sys.argv[2] if len(sys.argv) > 2 else "hist"
if mode == "hist":
    op = collections.Counter(); st = collections.Counter()
    for s_, e, sm in data:
        t = s_.split()
        o = t[1] if t[0].startswith("@") else t[0]
        o = o.split(".")[0]
        op[o] += e; st[o] += sm
    for o, e in op.most_common(40):
        print(f"{o:12s} {e:12d} {100*e/tot:5.1f}%  samples {st[o]}")
else:
    thr = float(mode) * mx
    for i, (s_, e, sm) in enumerate(data):
        if e >= thr:
            print(f"{i:5d} {e:10d} {sm:6d}  {s_}")
