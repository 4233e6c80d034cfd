#!/usr/bin/env python
"""Host->device link map of one box: H2D rate of every GPU alone, of every PAIR copying at once and of all GPUs at once
(pinned host memory, 256 MB per copy, CUDA events).  A pair whose concurrent rates drop well below the solo rates shares an
upstream link (PCIe switch / root port); a drop only with all GPUs active points at host memory or the VM's IO path.
Prints one JSON object.  usage: python scripts/h2d_linkmap.py [n_gpus]"""
import json
import sys
import time

import torch

n = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
MB = 256
host = [torch.empty(MB << 20, dtype=torch.uint8).pin_memory() for _ in range(n)]
dev = [torch.empty(MB << 20, dtype=torch.uint8, device=f"cuda:{i}") for i in range(n)]
streams = [torch.cuda.Stream(device=i) for i in range(n)]


def run(active, reps=4):
    ev = {}
    for i in active:
        with torch.cuda.device(i), torch.cuda.stream(streams[i]):
            dev[i].copy_(host[i], non_blocking=True)       # warm
    for i in active:
        torch.cuda.synchronize(i)
    for i in active:
        with torch.cuda.device(i), torch.cuda.stream(streams[i]):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                dev[i].copy_(host[i], non_blocking=True)
            e1.record()
            ev[i] = (e0, e1)
    out = {}
    for i in active:
        torch.cuda.synchronize(i)
        out[i] = reps * MB * (1 << 20) / (ev[i][0].elapsed_time(ev[i][1]) * 1e-3) / 1e9
    return out


solo = [run([i])[i] for i in range(n)]
pairs = {}
for i in range(n):
    for j in range(i + 1, n):
        r = run([i, j])
        pairs[f"{i}-{j}"] = [round(r[i], 1), round(r[j], 1)]
allr = run(list(range(n)))
print(json.dumps({"n_gpus": n, "copy_mb": MB, "solo_gbs": [round(v, 1) for v in solo], "pair_gbs": pairs,
                  "all_gbs": [round(allr[i], 1) for i in range(n)], "all_sum_gbs": round(sum(allr.values()), 1),
                  "when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime())}))
