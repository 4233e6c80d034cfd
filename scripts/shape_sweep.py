#!/usr/bin/env python
"""Parity sweep over odd image shapes: every stage against cv2, and 4-frame sequences against the reference restatement.
usage: python scripts/shape_sweep.py [n_random]"""
import os, sys
import numpy as np, cv2
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mobile_slam_b200 import BatchTracker, VS_RANSAC_OFF, config_from_synth
from mobile_slam_b200.synth import SynthStream
from oracle import camera as ocam, cv_reference as R


def shape_sweep(n_random=12, verbose=True):


    BASE = dict(max_cnt=60, min_dist=12, max_level=3, fisheye=0, model="PINHOLE", dist=(-0.1, 0.01, 1e-4, -1e-4), focal_length=460.0)
    shapes = [(64, 64), (65, 64), (64, 65), (71, 67), (96, 64), (100, 75), (127, 97), (128, 128), (129, 130), (160, 90), (255, 63),
              (256, 64), (257, 66), (320, 200), (333, 251), (401, 99), (640, 70), (70, 640), (1000, 65), (1023, 129), (1920, 96)]
    rng = np.random.default_rng(1)
    for _ in range(n_random):
        shapes.append((int(rng.integers(64, 900)), int(rng.integers(64, 600))))
    bad_list = []
    for (W, H) in shapes:
        cfg = dict(BASE, width=W, height=H, intr=(0.6 * W, 0.6 * W, W / 2.0, H / 2.0))
        if (W, H) in ((70, 640), (640, 70)):
            cfg["fisheye"] = 1
        try:
            tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF, capacity=256)
            bt = BatchTracker(tc, cam, 1)
        except Exception as e:  # noqa: BLE001
            print(W, H, "create failed:", str(e)[:80])
            continue
        st = SynthStream(W * 7 + H, max(W, 200), max(H, 200))
        fr = [np.ascontiguousarray(st.frame(k)[:H, :W]) for k in range(4)]
        msg = []
        e0, e1 = (cv2.createCLAHE(3.0, (8, 8)).apply(f) for f in fr[:2])
        if not np.array_equal(bt.stage_clahe(fr[0]), e0):
            msg.append("clahe")
        _, pyr = cv2.buildOpticalFlowPyramid(e0, (21, 21), 3, withDerivatives=False)
        got = bt.stage_pyramid(e0)
        if len(got) != len(pyr) or any(not np.array_equal(a, b) for a, b in zip(got, pyr)):
            msg.append(f"pyramid({len(got)} vs {len(pyr)} levels)")
        if not cfg["fisheye"]:
            c = bt.stage_good_features(e0, np.zeros((0, 2), np.float32), 60)
            r = cv2.goodFeaturesToTrack(e0, 60, 0.01, 12)
            r = np.zeros((0, 2), np.float32) if r is None else r.reshape(-1, 2)
            if not np.array_equal(c, r):
                sg, sr = {tuple(p) for p in c.tolist()}, {tuple(p) for p in r.tolist()}
                msg.append(f"gftt(common {len(sg & sr)}/{len(sr)})")
            if len(r):
                pts = np.concatenate([r, r + 0.37, np.array([[0.0, 0.0], [W - 1.0, H - 1.0], [W - 1.5, 0.5], [-3.0, H / 2]], np.float32)]).astype(np.float32)
                p1, s1 = bt.stage_lk(e0, e1, pts)
                p2, s2, _ = cv2.calcOpticalFlowPyrLK(e0, e1, pts.reshape(-1, 1, 2), None, winSize=(21, 21), maxLevel=3,
                                                     criteria=(cv2.TERM_CRITERIA_COUNT | cv2.TERM_CRITERIA_EPS, 30, 0.01))
                p2, s2 = p2.reshape(-1, 2), s2.reshape(-1)
                ok = (s1 == 1) & (s2 == 1)
                if np.mean(s1 == s2) < 0.995 or (ok.any() and np.abs(p1[ok] - p2[ok]).max() > 0.01):
                    msg.append(f"lk(status {np.mean(s1 == s2):.3f}, dpos {np.abs(p1[ok] - p2[ok]).max() if ok.any() else 0:.4f})")
        bt.close()
        bt = BatchTracker(tc, cam, 1)
        ref = R.RefFeatureTracker(R.config_from_synth(cfg, use_ransac=False), ocam.from_config(cfg))
        for k, f in enumerate(fr):
            r = bt.track([f], [0.05 * k])[0]
            ref.detect_and_track(f, 0.05 * k)
            ref.update_all_ids()
            if r.ids.tolist() != list(ref.ids) or r.track_cnt.tolist() != list(ref.track_cnt):
                common = len(set(r.ids.tolist()) & set(ref.ids))
                msg.append(f"seq frame {k} (n {len(r.ids)} vs {len(ref.ids)}, common ids {common})")
                break
            if len(ref.ids) and np.abs(r.cur_pts - np.array(ref.cur_pts, np.float32).reshape(-1, 2)).max() > 0.01:
                msg.append(f"seq frame {k} positions")
                break
        bt.close()
        if msg:
            bad_list.append(((W, H), msg))
        if verbose:
            print(W, H, "OK" if not msg else "MISMATCH: " + "; ".join(msg))
    if verbose:
        print("shapes with mismatches:", len(bad_list), "of", len(shapes))
    return bad_list


if __name__ == "__main__":
    shape_sweep(int(sys.argv[1]) if len(sys.argv) > 1 else 12)
