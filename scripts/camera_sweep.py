#!/usr/bin/env python
"""liftProjective parity sweep: many calibrations per camera model, pixels inside and far outside the image,
device (stage_undistort) against the oracle's fp64 restatement of the reference's camera classes.
usage: python scripts/camera_sweep.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mobile_slam_b200 import BatchTracker, CameraParams, TrackerConfig, VS_RANSAC_OFF
from mobile_slam_b200 import VS_CAM_KANNALA_BRANDT, VS_CAM_MEI, VS_CAM_PINHOLE, VS_CAM_SCARAMUZZA
from oracle import camera as ocam


def camera_sweep(verbose=True):
    W, H = 640, 480
    rng = np.random.default_rng(11)
    g = np.stack(np.meshgrid(np.linspace(-40, W + 40, 37), np.linspace(-40, H + 40, 29)), -1).reshape(-1, 2)
    pts = np.concatenate([g, np.stack([rng.uniform(0, W, 800), rng.uniform(0, H, 800)], 1),
                          [[320.0, 240.0], [0.0, 0.0], [W - 1.0, H - 1.0]]]).astype(np.float32)
    tc = TrackerConfig(col=W, row=H, ransac_mode=VS_RANSAC_OFF)
    cases = []
    for d in [(0, 0, 0, 0), (-0.28, 0.07, 2e-4, 2e-5), (0.1, -0.05, 0, 0), (-0.4, 0.2, -1e-3, 1e-3), (0, 0, 1e-3, 0), (1e-9, 0, 0, 0)]:
        for f in (250.0, 460.0, 900.0):
            cases.append(("PINHOLE", VS_CAM_PINHOLE, ocam.PINHOLE, (f, f * 1.01, 322.0, 238.0), d, {}))
    # (only trailing zero coefficients: EquidistantCamera.cc:585-596 lowers the degree once per zero k wherever it sits, so a
    # zero before a non-zero k divides by a zero leading coefficient in the reference - not a usable calibration)
    for d in [(0, 0, 0, 0), (0.0034, 0.0007, -0.002, 0.0002), (-0.01, 0.05, -0.08, 0.04), (0.05, 0, 0, 0),
              (0.1, -0.02, 0, 0), (-0.2, 0.1, 0, 0), (-0.3, 0.0, 0.0, 0.0)]:
        for f in (190.0, 380.0):
            cases.append(("KB", VS_CAM_KANNALA_BRANDT, ocam.KANNALA_BRANDT, (f, f, 320.0, 240.0), d, {}))
    for xi in (0.5, 1.0, 1.2, 2.2):
        for d in [(0, 0, 0, 0), (-0.2, 0.05, 1e-3, -5e-4)]:
            cases.append(("MEI", VS_CAM_MEI, ocam.MEI, (700.0, 705.0, 320.0, 240.0), d, dict(xi=xi)))
    for poly in [(-180.0, 0.0, 2.0e-3, -3.0e-6, 1.0e-8), (-250.0, 0.0, 1.5e-3, 0.0, 0.0)]:
        cases.append(("SCARA", VS_CAM_SCARAMUZZA, ocam.SCARAMUZZA, (1.0, 1.0, 318.0, 242.0), (0, 0, 0, 0),
                      dict(poly=poly, affine=(1.001, 0.002, -0.001))))
    bad = []
    for name, vm, om, (fx, fy, cx, cy), d, kw in cases:
        cam = CameraParams(vm, W, H, fx, fy, cx, cy, tuple(float(v) for v in d), **kw)
        bt = BatchTracker(tc, cam, 1)
        got = bt.stage_undistort(pts)
        bt.close()
        ref = ocam.Camera(om, fx, fy, cx, cy, tuple(float(v) for v in d), **kw).undistort_f32(pts)
        fin = np.isfinite(ref).all(1) & np.isfinite(got).all(1)
        same_nan = np.array_equal(np.isfinite(ref), np.isfinite(got))
        ulp = np.abs(got[fin].view(np.int32).astype(np.int64) - ref[fin].view(np.int32).astype(np.int64))
        tol = 1 if name == "KB" else 0
        frac_exact = float(np.mean(ulp == 0)) if ulp.size else 1.0
        ok = same_nan and (ulp.max() if ulp.size else 0) <= tol and frac_exact > 0.99
        if not ok:
            bad.append((name, (fx, fy, cx, cy), d, kw, int(ulp.max()) if ulp.size else 0, frac_exact, same_nan))
        if verbose:
            print(name, fx, d, kw, "OK" if ok else f"MISMATCH max ulp {int(ulp.max()) if ulp.size else 0} exact {frac_exact:.4f} nan-pattern {same_nan}")
    if verbose:
        print("mismatching calibrations:", len(bad), "of", len(cases))
    return bad


if __name__ == "__main__":
    camera_sweep()
