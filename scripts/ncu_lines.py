#!/usr/bin/env python
"""Join an .ncu-rep's per-SASS-instruction stall samples with source lines (nvdisasm -g on the kernel's cubin).
usage: ncu_lines.py <rep> <object.o> <kernel-substring> [top]"""
import csv, io, re, subprocess, sys, collections, os, tempfile
rep, obj, kname = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
lines = out.splitlines()
start = [i for i, l in enumerate(lines) if l.startswith('"Address"')][0]
rows = list(csv.reader(io.StringIO("\n".join(lines[start:]))))
h = rows[0]
ie, ss, src = h.index("Instructions Executed"), h.index("# Samples"), h.index("Source")
data = [(r[src].strip(), int(r[ie]), int(r[ss])) for r in rows[1:] if len(r) > ie]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cub = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")]
dis = ""
for c in cub:
    dis += subprocess.run(["nvdisasm", "-g", "-c", c], capture_output=True, text=True).stdout
# walk the disassembly of the wanted kernel: track current line from "//## File "...", line N" markers
cur_line, in_k, seq = None, False, []
for l in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m:
        in_k = kname in m.group(1)
        continue
    if not in_k:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur_line = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(.*?);", l)
    if m:
        seq.append(cur_line)
n = min(len(seq), len(data))
print(f"sass in report {len(data)}, in disassembly {len(seq)}")
agg_s, agg_e = collections.Counter(), collections.Counter()
for i in range(n):
    agg_s[seq[i]] += data[i][2]
    agg_e[seq[i]] += data[i][1]
ts, te = sum(agg_s.values()), sum(agg_e.values())
for k, v in agg_s.most_common(top):
    print(f"{str(k):32s} samples {v:7d} {100*v/max(ts,1):5.1f}%   exec {agg_e[k]:11d} {100*agg_e[k]/max(te,1):5.1f}%")
