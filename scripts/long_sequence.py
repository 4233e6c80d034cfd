#!/usr/bin/env python
"""300-frame free-running parity (the length of the reference's own trajectory parity test,
tests/test_vio_engine_parity.cpp:305) against the reference restatement, no RANSAC.
usage: python scripts/long_sequence.py [C1|C2|C3] [motion] [frames]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mobile_slam_b200 import BatchTracker, VS_RANSAC_OFF, config_from_synth
from mobile_slam_b200.synth import CONFIGS, SynthStream
from oracle import camera as ocam, cv_reference as R


def long_sequence(name="C2", motion=3.0, n_frames=300, verbose=True):
    cfg = CONFIGS[name]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    bt = BatchTracker(tc, cam, 1)
    ref = R.RefFeatureTracker(R.config_from_synth(cfg, use_ransac=False), ocam.from_config(cfg))
    st = SynthStream(91, cfg["width"], cfg["height"], motion=motion)
    exact = nref = ncommon = npos = 0
    dmax = 0.0
    first_div = None
    new_total = 0
    for k in range(n_frames):
        f = st.frame(k)
        r = bt.track([f], [0.05 * k])[0]
        ref.detect_and_track(f, 0.05 * k)
        ref.update_all_ids()
        same = r.ids.tolist() == list(ref.ids) and r.track_cnt.tolist() == list(ref.track_cnt)
        exact += same
        if not same and first_div is None:
            first_div = k
        gi = {i: j for j, i in enumerate(r.ids.tolist())}
        rp = np.array(ref.cur_pts, np.float32).reshape(-1, 2)
        nref += len(ref.ids)
        new_total += r.counters["n_new"] if "n_new" in r.counters else 0
        for j, i in enumerate(ref.ids):
            if i in gi and first_div is None:
                ncommon += 1
                dmax = max(dmax, float(np.abs(r.cur_pts[gi[i]] - rp[j]).max()))
        # id-free comparison: after a divergence event (one borderline cvRound / status flip renumbers every later
        # feature) the two trackers still follow the same image points
        if len(rp) and len(r.cur_pts):
            d = np.abs(rp[:, None, :] - r.cur_pts[None, :, :]).max(2).min(1)
            npos += int(np.sum(d <= 0.02))
    bt.close()
    out = dict(config=name, motion=motion, frames=n_frames, exact_frames=exact, first_divergence=first_div,
               ids_common_before_divergence=ncommon, pts_ref=nref, pts_matched_by_position=npos,
               pos_match_frac=npos / max(nref, 1), max_dpos_before_divergence=dmax, last_id=int(max(ref.ids)) if ref.ids else -1)
    if verbose:
        print(out)
    return out


if __name__ == "__main__":
    long_sequence(sys.argv[1] if len(sys.argv) > 1 else "C2", float(sys.argv[2]) if len(sys.argv) > 2 else 3.0,
                  int(sys.argv[3]) if len(sys.argv) > 3 else 300)
