#!/usr/bin/env python
"""Generates tests/golden/*.npz — the committed golden vectors of the hot path.

The reference (chansoopark98/Mobile-SLAM, src/frontend/feature_tracker.cpp) holds no golden vectors for
its feature tracker and cannot be compiled here (OpenCV C++ / Eigen / Ceres absent), so
the vectors are produced by the reference's own arithmetic dependency — OpenCV, through the cv2 wheel in
this image — on seeded synthetic frames:

  stages_small.npz   per-stage inputs/outputs of the five OpenCV calls feature_tracker.cpp makes
                     (CLAHE :102-106, buildOpticalFlowPyramid :123, calcOpticalFlowPyrLK :130-136,
                     goodFeaturesToTrack :185-186, circle :56,84) on a 240x208 frame pair
  sequence_small.npz the public vectors (ids, track_cnt, cur_pts, cur_undistorted_pts, pts_velocity) of
                     oracle/cv_reference.RefFeatureTracker (feature_tracker.cpp restated line by line, driving
                     cv2) over 8 frames of a 320x240 stream, RANSAC off, pinhole radtan camera
  camera.npz         liftProjective known-answer vectors: cv2.undistortPoints / cv2.fisheye.undistortPoints
                     (independent implementations of the same models, agree to ~1e-7) next to the float64
                     restatement oracle/camera.py

Run from the repo root:  python tests/golden/make_golden.py      (needs cv2; never run on the GPU box)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import cv2  # noqa: E402

from mobile_slam_b200.synth import CONFIGS, SynthStream  # noqa: E402
from oracle import camera as ocam  # noqa: E402
from oracle import cv_reference as R  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
META = dict(cv2_version=cv2.__version__, numpy=np.__version__)

SEQ_CFG = dict(width=320, height=240, max_cnt=60, min_dist=20, max_level=3, fisheye=0, model="PINHOLE",
               intr=(300.0, 300.0, 160.0, 120.0), dist=(-0.1, 0.01, 1e-4, -1e-4), focal_length=460.0)
SEQ_FISH_CFG = dict(width=256, height=256, max_cnt=50, min_dist=15, max_level=2, fisheye=1, model="KANNALA_BRANDT",
                    intr=(95.5, 95.5, 127.4, 128.6), dist=CONFIGS["C1"]["dist"], focal_length=460.0)


def stages():
    W, H = 240, 208
    st = SynthStream(42, W, H)
    f0, f1 = st.frame(3), st.frame(4)
    clahe = cv2.createCLAHE(3.0, (8, 8))
    e0, e1 = clahe.apply(f0), clahe.apply(f1)
    out = dict(frame0=f0, frame1=f1, clahe0=e0, clahe1=e1)
    # pyramid (REFLECT_101 padded by win inside cv2; we store the unpadded levels) + Scharr derivative of each level
    _, pyr = cv2.buildOpticalFlowPyramid(e0, (21, 21), 3, withDerivatives=True)
    n_levels = len(pyr) // 2
    out["n_levels"] = np.int32(n_levels)
    for l in range(n_levels):
        out[f"pyr{l}"] = np.ascontiguousarray(pyr[2 * l])
        out[f"deriv{l}"] = np.ascontiguousarray(pyr[2 * l + 1])
    # Shi-Tomasi: unmasked and masked by discs
    corners = cv2.goodFeaturesToTrack(e0, 40, 0.01, 10).reshape(-1, 2)
    out["gftt_all"] = corners
    mask = np.full((H, W), 255, np.uint8)
    kept = corners[::3].copy()
    for p in kept:
        cv2.circle(mask, (int(np.rint(p[0])), int(np.rint(p[1]))), 10, 0, -1)
    out["gftt_kept"] = kept
    out["gftt_mask"] = mask
    out["gftt_masked"] = cv2.goodFeaturesToTrack(e0, 25, 0.01, 10, mask=mask).reshape(-1, 2)
    out["min_eig"] = cv2.cornerMinEigenVal(e0, 3, ksize=3)
    # LK forward on the detected corners plus border / flat / out-of-frame probes
    extra = np.array([[0.4, 0.3], [W - 1.2, H - 1.6], [3.5, H / 2], [W / 2 + 0.25, 2.75], [W - 3.0, 5.0]], np.float32)
    pts = np.concatenate([corners, extra]).astype(np.float32)
    crit = (cv2.TERM_CRITERIA_COUNT | cv2.TERM_CRITERIA_EPS, 30, 0.01)
    nxt, status, _ = cv2.calcOpticalFlowPyrLK(e0, e1, pts, None, winSize=(21, 21), maxLevel=3, criteria=crit)
    out["lk_pts"], out["lk_next"], out["lk_status"] = pts, nxt.reshape(-1, 2), status.reshape(-1)
    # filled circles (cv::circle, thickness -1): row half-widths via rendering
    for r in (1, 2, 3, 5, 10, 20, 30, 57):
        m = np.zeros((2 * r + 5, 2 * r + 5), np.uint8)
        cv2.circle(m, (r + 2, r + 2), r, 255, -1)
        out[f"circle{r}"] = m
    np.savez_compressed(os.path.join(OUT, "stages_small.npz"), **out, **{f"meta_{k}": v for k, v in META.items()})


def sequence(name, cfg, sid, n_frames):
    st = SynthStream(sid, cfg["width"], cfg["height"], motion=2.0)
    ref = R.RefFeatureTracker(R.config_from_synth(cfg, use_ransac=False), ocam.from_config(cfg))
    out = dict(n_frames=np.int32(n_frames), stream_id=np.int32(sid), motion=np.float32(2.0))
    for k in range(n_frames):
        f = st.frame(k)
        ref.detect_and_track(f, 0.05 * k)
        ref.update_all_ids()
        out[f"frame{k}"] = f
        out[f"ids{k}"] = np.array(ref.ids, np.int32)
        out[f"cnt{k}"] = np.array(ref.track_cnt, np.int32)
        out[f"pts{k}"] = np.asarray(ref.cur_pts, np.float32).reshape(-1, 2)
        out[f"un{k}"] = np.asarray(ref.cur_undistorted_pts, np.float32).reshape(-1, 2)
        out[f"vel{k}"] = np.asarray(ref.pts_velocity, np.float32).reshape(-1, 2)
    np.savez_compressed(os.path.join(OUT, name), **out, **{f"meta_{k}": v for k, v in META.items()})


def camera():
    rng = np.random.default_rng(5)
    out = {}
    # pinhole radtan (EuRoC, config/euroc.yaml:18-27): cv2.undistortPoints iterates the same fixed point
    c2 = CONFIGS["C2"]
    fx, fy, cx, cy = c2["intr"]
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1.0]])
    D = np.array(c2["dist"], np.float64)
    px = np.stack([rng.uniform(0, c2["width"], 64), rng.uniform(0, c2["height"], 64)], 1)
    out["pinhole_px"] = px
    out["pinhole_cv"] = cv2.undistortPointsIter(px.reshape(-1, 1, 2), K, D, None, None,
                                                (cv2.TERM_CRITERIA_COUNT | cv2.TERM_CRITERIA_EPS, 200, 1e-14)).reshape(-1, 2)
    out["pinhole_oracle"] = ocam.from_config(c2).lift_batch(px)
    # Kannala-Brandt (TUM-VI, config/tum_vi_room1.yaml:18-26): cv2.fisheye uses k1..k4 = our k2..k5
    c1 = CONFIGS["C1"]
    fx, fy, cx, cy = c1["intr"]
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1.0]])
    D = np.array(c1["dist"], np.float64)
    ang = rng.uniform(0, 2 * np.pi, 64)
    rad = rng.uniform(0, 240, 64)
    px = np.stack([cx + rad * np.cos(ang), cy + rad * np.sin(ang)], 1)
    out["kb_px"] = px
    out["kb_cv"] = cv2.fisheye.undistortPoints(px.reshape(-1, 1, 2), K, D).reshape(-1, 2)
    out["kb_oracle"] = ocam.from_config(c1).lift_batch(px)
    np.savez_compressed(os.path.join(OUT, "camera.npz"), **out, **{f"meta_{k}": v for k, v in META.items()})


CLAHE_PAD_SHAPES = [(120, 90), (104, 100), (90, 120), (97, 96)]     # (W, H): one or both sides off the 8 x 8 tile grid


def clahe_padding():
    """cv::CLAHE on shapes that need its padding rule (a side that is a multiple of 8 still grows by 8)."""
    out = {}
    for w, h in CLAHE_PAD_SHAPES:
        img = cv2.GaussianBlur(np.random.default_rng(w * 1000 + h).integers(0, 256, (h, w), dtype=np.uint8), (0, 0), 2.0)
        out[f"in_{w}x{h}"] = img
        out[f"out_{w}x{h}"] = cv2.createCLAHE(3.0, (8, 8)).apply(img)
    np.savez_compressed(os.path.join(OUT, "clahe_padding.npz"), **out, **{f"meta_{k}": v for k, v in META.items()})


if __name__ == "__main__":
    if "--only-clahe-padding" in sys.argv:
        clahe_padding()
        raise SystemExit(0)
    stages()
    clahe_padding()
    sequence("sequence_small.npz", SEQ_CFG, 7, 8)
    sequence("sequence_fisheye.npz", SEQ_FISH_CFG, 9, 6)
    camera()
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))
