"""CUDA tracker vs the REFERENCE's own compiled code (oracle/_ref/libvsref.so — feature_tracker.cpp whole + the camera
models' liftProjective, built here from /root/reference by `make -C oracle ref`; the prebuilt .so travels to the GPU
box).  tests/test_oracle_ref.py shows the Python oracle bit-equal to this library on the CPU; here the product is
compared with it directly, so the parity claim for the glue and camera rows does not rest on the builder's Python.
"""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from mobile_slam_b200 import (BatchTracker, CameraParams, FeatureTracker, TrackerConfig, VS_CAM_MEI,  # noqa: E402
                              VS_CAM_SCARAMUZZA, VS_RANSAC_CALLBACK, VS_RANSAC_OFF, config_from_synth)
from mobile_slam_b200.synth import CONFIGS, SynthStream  # noqa: E402
from oracle import camera as ocam  # noqa: E402
from oracle import cv_reference as R  # noqa: E402
from oracle import ref_binding as RB  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref_lib():
    if not RB.available():
        pytest.skip("oracle/_ref/libvsref.so not shipped (build it with `make -C oracle ref` where /root/reference exists)")
    return RB.load()


def _f32_ref_undistort(t, pts):
    P = t.lift_batch(pts.astype(np.float64))
    return np.stack([P[:, 0] / P[:, 2], P[:, 1] / P[:, 2]], 1).astype(np.float32)


@pytest.mark.parametrize("name", ["C1", "C2", "C3", "C4"])
def test_undistort_vs_compiled_reference(built_lib, ref_lib, name):
    """vs_stage_undistort == (float)(P.x/P.z) of the compiled Camera::liftProjective: identical float32 for the
    closed-form models; <= 1 ulp for Kannala-Brandt (Newton on the device, companion-matrix eigenvalues there)."""
    c = CONFIGS[name]
    rng = np.random.default_rng(15)
    n = 10000 if c["model"] == "PINHOLE" else 3000
    pts = np.stack([rng.uniform(0, c["width"], n), rng.uniform(0, c["height"], n)], 1).astype(np.float32)
    tc, cam = config_from_synth(c, ransac_mode=VS_RANSAC_OFF)
    bt = BatchTracker(tc, cam, 1)
    got = bt.stage_undistort(pts)
    t = RB.CompiledRefTracker(R.config_from_synth(c), ocam.from_config(c))
    want = _f32_ref_undistort(t, pts)
    if c["model"] == "PINHOLE":
        assert np.array_equal(got, want)
    else:
        ulp = np.abs(got.view(np.int32) - want.view(np.int32))
        assert ulp.max() <= 1 and np.mean(ulp == 0) > 0.99, (ulp.max(), np.mean(ulp == 0))
    bt.close()
    t.close()


def test_undistort_mei_scaramuzza_vs_compiled_reference(built_lib, ref_lib):
    rng = np.random.default_rng(16)
    pts = np.stack([rng.uniform(0, 640, 10000), rng.uniform(0, 480, 10000)], 1).astype(np.float32)
    tc = TrackerConfig(col=640, row=480, ransac_mode=VS_RANSAC_OFF)
    rc = R.RefConfig(row=480, col=640)
    for xi in (1.0, 0.8, 1.9):
        d = (-0.2, 0.05, 1e-3, -5e-4)
        bt = BatchTracker(tc, CameraParams(VS_CAM_MEI, 640, 480, 700.0, 705.0, 320.0, 240.0, d, xi=xi), 1)
        t = RB.CompiledRefTracker(rc, ocam.Camera(ocam.MEI, 700.0, 705.0, 320.0, 240.0, d, xi=xi))
        got, want = bt.stage_undistort(pts), _f32_ref_undistort(t, pts)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), f"MEI xi={xi}"
        bt.close()
        t.close()
    poly, aff = (-180.0, 0.0, 2.0e-3, -3.0e-6, 1.0e-8), (1.001, 0.002, -0.001)
    bt = BatchTracker(tc, CameraParams(VS_CAM_SCARAMUZZA, 640, 480, 1.0, 1.0, 318.0, 242.0, poly=poly, affine=aff), 1)
    t = RB.CompiledRefTracker(rc, ocam.Camera(ocam.SCARAMUZZA, cx=318.0, cy=242.0, poly=poly, affine=aff))
    assert np.array_equal(bt.stage_undistort(pts), _f32_ref_undistort(t, pts))
    bt.close()
    t.close()


@pytest.mark.parametrize("name,n_frames,motion", [("C1", 30, 1.0), ("C2", 30, 1.0), ("C3", 12, 1.0)])
def test_sequence_vs_compiled_reference(built_lib, ref_lib, name, n_frames, motion):
    """Free-running sequence, RANSAC off on both sides: the CUDA tracker against the compiled reference tracker."""
    cfg = CONFIGS[name]
    st = SynthStream(41, cfg["width"], cfg["height"], motion=motion)
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    ft = FeatureTracker(tc)
    ft.setCamera(cam)
    ref = RB.CompiledRefTracker(R.config_from_synth(cfg, use_ransac=False), ocam.from_config(cfg))
    dt = 0.05
    worst = dict(pos=0.0, un=0.0, vel=0.0)
    exact = 0
    for k in range(n_frames):
        f = st.frame(k)
        ft.detectAndTrack(f, dt * k)
        i = 0
        while ft.updateID(i):
            i += 1
        ref.detect_and_track(f, dt * k)
        ref.update_all_ids()
        if ft.ids == ref.ids and ft.track_cnt == ref.track_cnt:
            exact += 1
            worst["pos"] = max(worst["pos"], float(np.abs(ft.cur_pts - ref.cur_pts).max()))
            worst["un"] = max(worst["un"], float(np.abs(ft.cur_undistorted_pts - ref.cur_undistorted_pts).max()))
            worst["vel"] = max(worst["vel"], float(np.abs(ft.pts_velocity - ref.pts_velocity).max()))
    assert exact >= int(0.9 * n_frames), f"{exact}/{n_frames} frames identical in ids + track_cnt"
    assert worst["pos"] <= 0.01, worst
    assert worst["un"] <= 1e-4, worst
    assert worst["vel"] <= 2 * 1e-4 / dt, worst     # what the 1e-4 undistortion bound implies for a difference / dt
    ref.close()


def test_sequence_with_cv_ransac_vs_compiled_reference(built_lib, ref_lib):
    """rejectWithFundamentalMatrix: callback mode (cv2.findFundamentalMat on the device's lifted points) against the
    compiled reference calling the same cv2 function — incl. the compiled edge-recovery loop (edge factor > 0)."""
    cfg = CONFIGS["C2"]
    n_frames = 24
    st = SynthStream(43, cfg["width"], cfg["height"], motion=3.0)
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_CALLBACK)
    bt = BatchTracker(tc, cam, 1)

    def reject(stream, a, b, px):
        F, status = cv2.findFundamentalMat(a, b, cv2.FM_RANSAC, tc.f_threshold, 0.99)
        return np.ones(len(a), np.uint8) if status is None else status.reshape(-1)
    bt.set_reject_callback(reject)
    ref = RB.CompiledRefTracker(R.config_from_synth(cfg), ocam.from_config(cfg))
    exact = ran = 0
    for k in range(n_frames):
        f = st.frame(k)
        r = bt.track([f], [0.05 * k])[0]
        ref.detect_and_track(f, 0.05 * k)
        ref.update_all_ids()
        ran += bool(ref.stage.get("ransac_ran"))
        if r.ids.tolist() == ref.ids and r.track_cnt.tolist() == ref.track_cnt:
            exact += 1
            assert np.abs(r.cur_pts - ref.cur_pts).max() <= 0.01
            assert np.abs(r.cur_undistorted_pts - ref.cur_undistorted_pts).max() <= 1e-4
    assert ran >= n_frames // 2
    assert exact >= int(0.85 * n_frames), f"{exact}/{n_frames} frames identical"
    bt.close()
    ref.close()


def test_frontend_flags_mid_sequence_vs_compiled_reference(built_lib, ref_lib):
    """g_config.feature_tracker.equalize / .fisheye and camera.focal_length are re-read by the reference inside every call
    (feature_tracker.cpp:39-60,102,214-222): toggled between frames (vs_tracker_set_frontend_flags through the mirror's
    per-call config), the CUDA tracker must keep following the compiled reference without losing its tracking state."""
    cfg = dict(CONFIGS["C2"], width=376, height=240, max_cnt=100, min_dist=20, intr=(230.0, 230.0, 188.0, 120.0))
    st = SynthStream(9, 376, 240, motion=1.5)
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    ft = FeatureTracker(tc)
    ft.setCamera(cam)
    rc = R.config_from_synth(cfg, use_ransac=False)
    ref = RB.CompiledRefTracker(rc, ocam.from_config(cfg))
    exact = 0
    n_frames = 14
    for k in range(n_frames):
        if k == 3:
            rc.fisheye = ft.config.fisheye = 1
        if k == 6:
            rc.equalize = ft.config.equalize = 0
        if k == 9:
            rc.fisheye = ft.config.fisheye = 0
            rc.equalize = ft.config.equalize = 1
        f = st.frame(k)
        ft.detectAndTrack(f, 0.05 * k)
        i = 0
        while ft.updateID(i):
            i += 1
        ref.detect_and_track(f, 0.05 * k)
        ref.update_all_ids()
        assert len(ft.ids) > 40, f"frame {k}: tracking state lost ({len(ft.ids)} features)"
        if ft.ids == ref.ids and ft.track_cnt == ref.track_cnt:
            exact += 1
            assert np.abs(ft.cur_pts - ref.cur_pts).max() <= 0.01
    assert exact >= n_frames - 2, f"{exact}/{n_frames} frames identical in ids + track_cnt"
    ref.close()
