#!/usr/bin/env python
"""Device RANSAC vs cv2.findFundamentalMat over many synthetic two-view sets: kept-inlier / rejected-outlier rates."""
import sys, numpy as np, cv2
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from test_gpu_stages import _two_view
from test_gpu_stages import _cfg, config_from_synth, BatchTracker, VS_RANSAC_OFF
tc, cam = config_from_synth(_cfg("C2"), ransac_mode=VS_RANSAC_OFF, capacity=512)
t = BatchTracker(tc, cam, 1)
for n, frac in [(150, 0.2), (150, 0.4), (150, 0.5), (60, 0.1), (400, 0.3)]:
    r = []
    for seed in range(30):
        a, b, outl = _two_view(n, frac, 100 + seed)
        st = t.stage_ransac(a, b).astype(bool)
        cvm = cv2.findFundamentalMat(a, b, cv2.FM_RANSAC, 1.0, 0.99)[1].reshape(-1).astype(bool)
        inl = ~outl
        r.append((st[inl].mean(), cvm[inl].mean(), (~st[outl]).mean() if outl.any() else 1, (~cvm[outl]).mean() if outl.any() else 1, (st == cvm).mean()))
    r = np.array(r)
    print(n, frac, "kept dev %.3f cv %.3f | min dev %.3f cv %.3f | rej dev %.3f cv %.3f | agree mean %.3f min %.3f" % (
        r[:, 0].mean(), r[:, 1].mean(), r[:, 0].min(), r[:, 1].min(), r[:, 2].mean(), r[:, 3].mean(), r[:, 4].mean(), r[:, 4].min()))
