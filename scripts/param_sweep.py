#!/usr/bin/env python
"""Parity sweep over tracker parameters: 6-frame sequences against the reference restatement (no RANSAC), one
parameter combination at a time.  usage: python scripts/param_sweep.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mobile_slam_b200 import BatchTracker, VS_RANSAC_OFF, config_from_synth
from mobile_slam_b200.synth import CONFIGS, SynthStream
from oracle import camera as ocam, cv_reference as R


def param_sweep(n_random=12, verbose=True):


    cases = []
    for win in (5, 7, 9, 11, 13, 15, 17, 19, 23, 25, 27, 29, 31):
        cases.append(("C2", dict(lk_window_size=win)))
    for lv in (0, 1, 2, 4, 5):
        cases.append(("C2", dict(lk_pyramid_levels=lv)))
    for md in (1, 2, 3, 5, 8, 50, 100, 200):
        cases.append(("C2", dict(min_dist=md)))
    for mc in (1, 2, 10, 31, 300, 500, 1000):
        cases.append(("C2", dict(max_cnt=mc, min_dist=10)))
    for it, eps in ((1, 0.01), (5, 0.1), (20, 0.03), (100, 0.001), (30, 0.0)):
        cases.append(("C2", dict(lk_iterations=it, lk_eps=eps)))
    cases += [("C1", dict(equalize=0)), ("C2", dict(equalize=0)), ("C1", dict(fisheye=0)), ("C2", dict(fisheye=1)),
              ("C3", dict(lk_window_size=15)), ("C1", dict(min_dist=7, max_cnt=400)), ("C3", dict(max_cnt=500))]
    bad_list = []
    for name, over in cases:
        cfg = dict(CONFIGS[name])
        for k in ("min_dist", "max_cnt", "fisheye"):
            if k in over:
                cfg[k] = over[k]
        extra = {k: v for k, v in over.items() if k not in ("min_dist", "max_cnt", "fisheye")}
        if "lk_pyramid_levels" in extra:
            cfg["max_level"] = extra.pop("lk_pyramid_levels")
        try:
            tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF, capacity=max(cfg["max_cnt"] + 64, 256), **extra)
            bt = BatchTracker(tc, cam, 1)
        except Exception as e:  # noqa: BLE001
            print(name, over, "create failed:", str(e)[:100])
            continue
        ref = R.RefFeatureTracker(R.config_from_synth(cfg, use_ransac=False, **extra), ocam.from_config(cfg))
        st = SynthStream(17, cfg["width"], cfg["height"], motion=1.5)
        msg, exact, nref, ncommon, dmax = "", 0, 0, 0, 0.0
        for k in range(6):
            f = st.frame(k)
            r = bt.track([f], [0.05 * k])[0]
            ref.detect_and_track(f, 0.05 * k)
            ref.update_all_ids()
            exact += r.ids.tolist() == list(ref.ids) and r.track_cnt.tolist() == list(ref.track_cnt)
            gi = {i: j for j, i in enumerate(r.ids.tolist())}
            rp = np.array(ref.cur_pts, np.float32).reshape(-1, 2)
            nref += len(ref.ids)
            for j, i in enumerate(ref.ids):
                if i in gi:
                    ncommon += 1
                    dmax = max(dmax, float(np.abs(r.cur_pts[gi[i]] - rp[j]).max()))
        bt.close()
        # a single pyramid level cannot absorb this motion: LK stops at the iteration cap without converging and the
        # float rounding differences of 30 un-damped steps show (0.02 px observed); every converged case is < 0.01 px
        tol = 0.03 if cfg["max_level"] == 0 else 0.01
        ok = exact >= 5 and ncommon >= 0.99 * nref and dmax <= tol
        if not ok:
            bad_list.append((name, over, exact, ncommon, nref, dmax))
        if verbose:
            print(name, over, "OK" if ok else "MISMATCH", f"exact frames {exact}/6, ids common {ncommon}/{nref}, max dpos {dmax:.4f}")
    if verbose:
        print("cases with mismatches:", len(bad_list), "of", len(cases))
    return bad_list


if __name__ == "__main__":
    param_sweep()
