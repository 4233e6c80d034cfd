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
# feature bank: a long estimator loop fed from the tracker on the device (window slides every frame), then create / destroy
from mobile_slam_b200.feature_bank import OP_NONE, OP_REMOVE_BACK, OP_REMOVE_FRONT, WINDOW_SIZE, FeatureBank
bt = BatchTracker(tc, cam, S)
bank = FeatureBank(S, capacity=1024)
free4 = torch.cuda.mem_get_info()[0]
fcs = np.zeros(S, np.int32)
sizes, n_bank = [], min(n_batches, 1500)
for i in range(n_bank):
    bt.track_pinned(pinned[idx(i)].data_ptr(), H * W, (C.c_double * S)(*([0.05 * i] * S)))
    keys = bank.add_frames_from_tracker(bt, fcs)
    full = fcs >= WINDOW_SIZE
    bank.slide(np.where(full, np.where(keys, OP_REMOVE_BACK, OP_REMOVE_FRONT), OP_NONE).astype(np.int32), fcs)
    fcs = np.where(full, fcs, fcs + 1).astype(np.int32)
    if i % 250 == 249:
        sizes.append(int(np.mean([len(bank.export(s_)[0]) for s_ in range(0, S, 4)])))
free5 = torch.cuda.mem_get_info()[0]
print(f"feature bank: {n_bank} frames x {S} streams, features per bank {min(sizes)}..{max(sizes)}, growth {(free4 - free5) / 2**20:.1f} MiB")
assert max(sizes) < 1024 and min(sizes) > 60 and abs(free4 - free5) < 8 * 2**20
bank.close(); bt.close()
for k in range(60):
    b3 = FeatureBank(8, capacity=512)
    b3.add_frames(0, [{i: [0.01 * i, 0.0, 1.0, 0, 0, 0, 0] for i in range(50)}] * 8)
    b3.close()
free6 = torch.cuda.mem_get_info()[0]
print(f"after 60 bank create/add/destroy cycles: {(free0 - free6) / 2**20:.1f} MiB not returned")
assert free0 - free6 < 64 * 2**20
print("soak ok")
