"""The TMA variant of the LK tile staging (VS_LK_TMA=1: cp.async.bulk.tensor + mbarrier instead of cp.async, csrc/lk.cu) must
track bit-identically to the shipped kernel.  The switch is read once per process, so each arm runs in its own interpreter."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, %r)
from mobile_slam_b200 import BatchTracker, VS_RANSAC_OFF, config_from_synth
from mobile_slam_b200.synth import CONFIGS, SynthStream
out = {}
for name, B in (("C2", 3), ("C3", 20)):
    cfg = CONFIGS[name]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    tc.fb_check = 1 if name == "C2" else 0
    bt = BatchTracker(tc, cam, B)
    sts = [SynthStream(50 + s, cfg["width"], cfg["height"], motion=2.0) for s in range(B)]
    for k in range(6):
        res = bt.track([st.frame(k) for st in sts], [0.05 * k] * B)
    for s in (0, B - 1):
        out[f"{name}_{s}_ids"] = res[s].ids
        out[f"{name}_{s}_cnt"] = res[s].track_cnt
        out[f"{name}_{s}_pts"] = res[s].cur_pts
        out[f"{name}_{s}_vel"] = res[s].pts_velocity
    bt.close()
np.savez(sys.argv[1], **out)
"""


def test_lk_tma_staging_tracks_bit_identically(tmp_path, built_lib):
    outs = []
    for tma in ("0", "1"):
        p = tmp_path / f"tma{tma}.npz"
        env = dict(os.environ, VS_LK_TMA=tma)
        r = subprocess.run([sys.executable, "-c", SCRIPT % ROOT, str(p)], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(np.load(p))
    a, b = outs
    assert sorted(a.files) == sorted(b.files) and len(a.files) == 16
    for k in a.files:
        assert a[k].shape == b[k].shape and np.array_equal(a[k].view(np.uint8), b[k].view(np.uint8)), k
    assert len(a["C2_0_ids"]) > 100
