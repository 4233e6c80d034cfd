"""Single-stream latency probe: N frames of one synthetic stream through vs_tracker_track (host frame in, results out).
usage: python scripts/latency_probe.py [C1|C2] [frames] [off]   (off = without the F-matrix RANSAC)"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mobile_slam_b200 import BatchTracker, VS_RANSAC_OFF, VS_RANSAC_DEVICE, config_from_synth
from mobile_slam_b200.synth import CONFIGS, SynthStream
import ctypes as C
name = sys.argv[1] if len(sys.argv) > 1 else "C1"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 60
cfg = CONFIGS[name]
st = SynthStream(77, cfg["width"], cfg["height"])
frames = [st.frame(k) for k in range(24)]
mode = VS_RANSAC_OFF if (len(sys.argv) > 3 and sys.argv[3] == 'off') else VS_RANSAC_DEVICE
tc, cam = config_from_synth(cfg, ransac_mode=mode)
bt = BatchTracker(tc, cam, 1)
ts, dev = [], []
for i in range(n):
    k = i % 46
    f = frames[k if k < 24 else 46 - k]
    t0 = time.perf_counter()
    bt.track([f], [0.05 * i], collect=False)
    ts.append(time.perf_counter() - t0)
    dev.append(bt.last_device_ms)
ts = np.array(ts[5:]) * 1e6
print(f"{name}: p50 {np.percentile(ts,50):.1f} us  p95 {np.percentile(ts,95):.1f} us  device p50 {np.percentile(dev[5:],50)*1e3:.1f} us")
bt.set_profiling(True)
for i in range(n, n + 40):
    k = i % 46
    bt.track([frames[k if k < 24 else 46 - k]], [0.05 * i], collect=False)
print({k: round(1e3 * ms / max(c, 1), 1) for k, (ms, c) in bt.profile().items() if c})
