#!/usr/bin/env python
"""Soak: thousands of batches through submit/wait (and a handle create/destroy loop), watching for errors, hangs,
feature-count collapse and device-memory growth.  usage: python scripts/soak.py [batches] [streams]"""
import os, sys, time, ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mobile_slam_b200 import BatchTracker, VS_RANSAC_DEVICE, config_from_synth
from mobile_slam_b200.synth import CONFIGS, SynthStream

n_batches = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
S = int(sys.argv[2]) if len(sys.argv) > 2 else 16
cfg = CONFIGS["C1"]
W, H = cfg["width"], cfg["height"]
streams = [SynthStream(50 + s, W, H, motion=1.0 + s % 3) for s in range(S)]
n_buf = 40
frames = np.stack([np.stack([st.frame(k) for st in streams]) for k in range(n_buf)])   # [n_buf][S][H][W]
pinned = torch.from_numpy(frames).pin_memory()
free0 = torch.cuda.mem_get_info()[0]
tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_DEVICE, pipeline_lanes=2)
bt = BatchTracker(tc, cam, S)
free1 = torch.cuda.mem_get_info()[0]
def idx(i):
    k = i % (2 * n_buf - 2)
    return k if k < n_buf else 2 * n_buf - 2 - k          # bounce: continuous motion
t0 = time.time()
pub = []
bt.submit_pinned(pinned[idx(0)].data_ptr(), H * W, (C.c_double * S)(*([0.0] * S)))
for i in range(1, n_batches):
    bt.submit_pinned(pinned[idx(i)].data_ptr(), H * W, (C.c_double * S)(*([0.05 * i] * S)))
    r = bt.wait(collect=(i % 250 == 0))
    if r is not None:
        pub.append(float(np.mean([np.sum(x.track_cnt > 1) for x in r])))
bt.wait()
dt = time.time() - t0
free2 = torch.cuda.mem_get_info()[0]
bt.close()
print(f"{n_batches} batches x {S} streams in {dt:.1f} s = {n_batches * S / dt:.0f} frames/s; published per stream {min(pub):.0f}..{max(pub):.0f}")
print(f"device memory: handle {(free0 - free1) / 2**20:.0f} MiB, growth while running {(free1 - free2) / 2**20:.1f} MiB")
assert min(pub) > 60 and abs(free1 - free2) < 8 * 2**20
for k in range(60):                                        # create / destroy
    b2 = BatchTracker(tc, cam, 4)
    b2.track([frames[0][s] for s in range(4)], [0.0] * 4)
    b2.close()
free3 = torch.cuda.mem_get_info()[0]
print(f"after 60 create/track/destroy cycles: {(free0 - free3) / 2**20:.1f} MiB not returned")
assert free0 - free3 < 64 * 2**20
print("soak ok")
