"""Pins the oracle to code COMPILED FROM /root/reference (oracle/_ref/libvsref.so, recipe: `make -C oracle ref`).

libvsref.so = the reference's src/frontend/feature_tracker.cpp, whole and unmodified, + the liftProjective-path
functions of src/common/camera_models/*.cc, with OpenCV calls forwarded to cv2 (oracle/ref_binding.py).  Every
comparison below is BIT-EXACT: the Python restatements (oracle/cv_reference.py, oracle/camera.py) must reproduce the
compiled reference's ids, track counts, pixel positions, undistorted points, velocities and masks to the last bit —
except the Kannala-Brandt root, where Eigen's EigenSolver is stood in by LAPACK on both sides (bit-equal to each
other; the stand-in is named in oracle/ref_shim/Eigen/Dense).
"""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from mobile_slam_b200.synth import CONFIGS, SynthStream  # noqa: E402
from oracle import camera as ocam  # noqa: E402
from oracle import cv_reference as R  # noqa: E402
from oracle import ref_binding as RB  # noqa: E402


@pytest.fixture(scope="module")
def ref_lib():
    if not RB.available():
        import os
        import subprocess
        if os.path.exists("/root/reference/src/frontend/feature_tracker.cpp"):
            subprocess.check_call(["make", "-s", "-C", os.path.dirname(RB.LIB_PATH) + "/..", "ref"])
    if not RB.available():
        pytest.skip("oracle/_ref/libvsref.so not built and /root/reference absent")
    return RB.load()


CAMERAS = {
    "pinhole_radtan": ocam.Camera(ocam.PINHOLE, 461.6, 460.3, 363.0, 248.1, (-0.2917, 0.08228, 5.333e-05, -1.578e-04)),
    "pinhole_nodist": ocam.Camera(ocam.PINHOLE, 552.554261, 552.554261, 682.049453, 238.769549, (0.0, 0.0, 0.0, 0.0)),
    "pinhole_strong": ocam.Camera(ocam.PINHOLE, 300.0, 310.0, 320.5, 239.5, (-0.4, 0.2, 1e-3, -2e-3)),
    "mei": ocam.Camera(ocam.MEI, 700.0, 705.0, 376.0, 240.0, (-0.2, 0.1, 1e-4, -2e-4), xi=1.8),
    "mei_xi1": ocam.Camera(ocam.MEI, 400.0, 400.0, 320.0, 240.0, (-0.1, 0.02, 0.0, 1e-4), xi=1.0),
    "mei_nodist": ocam.Camera(ocam.MEI, 500.0, 500.0, 320.0, 240.0, (0.0, 0.0, 0.0, 0.0), xi=0.9),
    "scaramuzza": ocam.Camera(ocam.SCARAMUZZA, cx=376.2, cy=239.7, poly=(-260.0, 0.0, 1.9e-3, -2.0e-6, 1.1e-8),
                              affine=(1.0006, -2e-4, 3e-4)),
    "kb_tumvi": ocam.from_config(CONFIGS["C1"]),
    "kb_k2_only": ocam.Camera(ocam.KANNALA_BRANDT, 280.0, 281.0, 320.0, 240.0, (-0.013, 0.0, 0.0, 0.0)),
    "kb_k2k3": ocam.Camera(ocam.KANNALA_BRANDT, 280.0, 281.0, 320.0, 240.0, (-0.013, 0.02, 0.0, 0.0)),
    "kb_none": ocam.Camera(ocam.KANNALA_BRANDT, 280.0, 281.0, 320.0, 240.0, (0.0, 0.0, 0.0, 0.0)),
}


def _pixels(n, w, h, seed):
    rng = np.random.default_rng(seed)
    px = np.stack([rng.uniform(-8.0, w + 8.0, n), rng.uniform(-8.0, h + 8.0, n)], 1)
    px[:64, 0] = np.arange(64) * (w / 64.0)          # integer / half-integer coordinates too
    px[:64, 1] = np.arange(64) * (h / 64.0) + 0.5
    return px.astype(np.float32).astype(np.float64)   # the tracker lifts float32 pixel positions


@pytest.mark.parametrize("name", sorted(CAMERAS))
def test_lift_projective_bit_equal(ref_lib, name):
    """oracle/camera.py == the compiled Camera::liftProjective, fp64 bit for bit, on 10^4 pixels per model."""
    c = CAMERAS[name]
    cfg = R.RefConfig(row=480, col=752)
    t = RB.CompiledRefTracker(cfg, c)
    n = 10000 if c.model != ocam.KANNALA_BRANDT or name == "kb_none" else 2500   # KB: one LAPACK call per pixel
    px = _pixels(n, 752, 480, 5)
    got = t.lift_batch(px)
    want = np.array([c.lift_projective(float(u), float(v)) for u, v in px])
    same = (got.view(np.uint64) == want.view(np.uint64)) | (np.isnan(got) & np.isnan(want))
    assert same.all(), f"{name}: {np.count_nonzero(~same)} of {same.size} doubles differ; max |d| {np.nanmax(np.abs(got - want))}"
    # the vectorised path the Python tracker uses must agree with the scalar restatement as well
    vb = c.lift_batch(px)
    un_a = np.stack([vb[:, 0] / vb[:, 2], vb[:, 1] / vb[:, 2]], 1).astype(np.float32)
    un_b = np.stack([got[:, 0] / got[:, 2], got[:, 1] / got[:, 2]], 1).astype(np.float32)
    ok = (un_a == un_b) | (np.isnan(un_a) & np.isnan(un_b))
    assert ok.mean() >= (1.0 if c.model != ocam.KANNALA_BRANDT else 0.999), f"{name}: float32 undistorted points differ"
    t.close()


def test_in_border_matches_compiled(ref_lib):
    cfg = R.RefConfig(row=480, col=752)
    cam_ = CAMERAS["pinhole_nodist"]
    t = RB.CompiledRefTracker(cfg, cam_)
    py = R.RefFeatureTracker(cfg, cam_)
    rng = np.random.default_rng(0)
    pts = np.concatenate([
        np.stack([rng.uniform(-3, 755, 4000), rng.uniform(-3, 483, 4000)], 1),
        np.array([[0.5, 10], [1.5, 10], [2.5, 10], [749.5, 10], [750.5, 10], [751.5, 10], [10, 0.5], [10, 1.5],
                  [10, 477.5], [10, 478.5], [10, 479.5], [0.49999997, 0.50000006], [750.49994, 478.50003]])]).astype(np.float32)
    for p in pts:
        assert t.in_border(p) == py.in_border(p), p
    t.close()


def _run_pair(cfg, cam_, frames, times, rc_kw=None, mutate=None):
    rc_a = R.config_from_synth(cfg, **(rc_kw or {}))
    rc_b = R.config_from_synth(cfg, **(rc_kw or {}))
    a = R.RefFeatureTracker(rc_a, cam_)
    b = RB.CompiledRefTracker(rc_b, cam_)
    ran = 0
    for k, (f, t) in enumerate(zip(frames, times)):
        if mutate:
            mutate(k, rc_a)
            mutate(k, rc_b)
        a.detect_and_track(f, t)
        b.detect_and_track(f, t)
        # before updateID: new points carry id -1 on both sides
        assert a.ids == b.ids, f"frame {k}: ids differ before updateID"
        a.update_all_ids()
        b.update_all_ids()
        assert a.ids == b.ids, f"frame {k}: ids differ"
        assert a.track_cnt == b.track_cnt, f"frame {k}: track_cnt differ"
        for name in ("cur_pts", "cur_undistorted_pts", "pts_velocity"):
            x, y = getattr(a, name), getattr(b, name)
            assert x.shape == y.shape and np.array_equal(x.view(np.uint32), y.view(np.uint32)), \
                f"frame {k}: {name} not bit-equal (max |d| {np.abs(x - y).max() if x.shape == y.shape else 'shape'})"
        assert np.array_equal(a.mask, b.mask), f"frame {k}: setMask image differs"
        assert bool(a.stage.get("ransac_ran")) == bool(b.stage.get("ransac_ran")), f"frame {k}: RANSAC gate differs"
        assert a.published() == b.published()
        ran += bool(b.stage.get("ransac_ran"))
    b.close()
    return a, ran


@pytest.mark.parametrize("name,n_frames,motion", [("C1", 14, 3.0), ("C2", 14, 3.0), ("C3", 6, 2.0), ("C2", 10, 0.3)])
def test_python_oracle_equals_compiled_reference(ref_lib, name, n_frames, motion):
    """Free-running sequences incl. cv2.findFundamentalMat: the restated glue is bit-identical to the compiled one."""
    cfg = CONFIGS[name]
    st = SynthStream(11, cfg["width"], cfg["height"], motion=motion)
    frames = [st.frame(k) for k in range(n_frames)]
    _, ran = _run_pair(cfg, ocam.from_config(cfg), frames, [0.05 * k for k in range(n_frames)])
    if motion >= 2.0:
        assert ran >= n_frames // 2, "the RMS gate should let the F-matrix RANSAC run on fast motion"
    else:
        assert ran == 0, "slow motion: the RMS gate (< 2 px) must skip the RANSAC on both sides"


def test_edge_recovery_and_small_config(ref_lib):
    """f_threshold_edge_factor > 0 (feature_tracker.cpp:236-285) with a tight threshold so that points are rejected
    and some restored; 320x240 frames."""
    cfg = dict(CONFIGS["C2"], width=320, height=240, max_cnt=120, min_dist=12, intr=(300.0, 300.0, 160.0, 120.0),
               dist=(-0.25, 0.07, 1e-4, -1e-4))
    st = SynthStream(5, 320, 240, motion=3.0)
    frames = [st.frame(k) for k in range(12)]
    a, ran = _run_pair(cfg, ocam.from_config(cfg), frames, [0.05 * k for k in range(12)],
                       rc_kw=dict(f_threshold=0.15, f_threshold_edge_factor=4.0))
    assert ran >= 6


def test_degenerate_inputs(ref_lib):
    """featureless frames, a scene cut, repeated / decreasing timestamps, equalize off, fisheye mask."""
    cfg = dict(CONFIGS["C1"], width=256, height=256, max_cnt=80, min_dist=15,
               intr=(95.0, 95.0, 127.4, 128.6))
    s1, s2 = SynthStream(1, 256, 256, motion=2.0), SynthStream(2, 256, 256, motion=2.0)
    flat = np.full((256, 256), 90, np.uint8)
    frames = [s1.frame(0), s1.frame(1), flat, flat, s1.frame(2), s1.frame(3), s2.frame(0), s2.frame(1), s2.frame(2),
              s2.frame(3)]
    times = [0.0, 0.05, 0.10, 0.15, 0.20, 0.20, 0.25, 0.24, 0.30, 0.30005]
    _run_pair(cfg, ocam.from_config(cfg), frames, times)
    _run_pair(cfg, ocam.from_config(cfg), frames, times, rc_kw=dict(equalize=0))


def test_runtime_parameter_mutation(ref_lib):
    """g_config changes between frames (vio_engine.cpp:354-389 mutators): max_cnt, min_dist, f_threshold, LK criteria."""
    cfg = dict(CONFIGS["C2"], width=376, height=240, max_cnt=100, min_dist=20, intr=(230.0, 230.0, 188.0, 120.0))
    st = SynthStream(9, 376, 240, motion=2.5)
    frames = [st.frame(k) for k in range(12)]

    def mutate(k, rc):
        if k == 3:
            rc.max_cnt, rc.min_dist = 60, 25
        if k == 6:
            rc.lk_iterations, rc.lk_eps, rc.f_threshold = 20, 0.03, 0.5
        if k == 9:
            rc.max_cnt, rc.min_dist, rc.lk_window_size = 140, 10, 15
    _run_pair(cfg, ocam.from_config(cfg), frames, [0.05 * k for k in range(12)], mutate=mutate)

    def flags(k, rc):                   # the fields the reference re-reads inside every call besides the mutators'
        if k == 3:
            rc.fisheye = 1              # the fisheye mask is built lazily on the first frame that has the flag (ft:39-60)
        if k == 6:
            rc.equalize = 0
        if k == 8:
            rc.fisheye, rc.equalize, rc.focal_length = 0, 1, 300.0
    _run_pair(cfg, ocam.from_config(cfg), frames, [0.05 * k for k in range(12)], mutate=flags)


def test_other_camera_models_in_sequence(ref_lib):
    """Mei and Scaramuzza cameras through the whole compiled tracker (undistorted points + velocity)."""
    cfg = dict(CONFIGS["C2"], width=376, height=240, max_cnt=80, min_dist=20)
    st = SynthStream(4, 376, 240, motion=2.0)
    frames = [st.frame(k) for k in range(6)]
    mei = ocam.Camera(ocam.MEI, 350.0, 352.0, 188.0, 120.0, (-0.2, 0.1, 1e-4, -2e-4), xi=1.6)
    sca = ocam.Camera(ocam.SCARAMUZZA, cx=188.2, cy=119.7, poly=(-130.0, 0.0, 3.8e-3, -8.0e-6, 9e-8),
                      affine=(1.0006, -2e-4, 3e-4))
    for c in (mei, sca):
        _run_pair(cfg, c, frames, [0.05 * k for k in range(6)])


# ---- a13: readIntrinsicParameter (CameraFactory YAML route) of the compiled reference vs the product's readers ---------
_YAMLS = {
    "pinhole": """%YAML:1.0
---
model_type: PINHOLE
camera_name: camera
image_width: 752
image_height: 480
distortion_parameters:
   k1: -2.917e-01
   k2: 8.228e-02
   p1: 5.333e-05
   p2: -1.578e-04
projection_parameters:
   fx: 4.616e+02
   fy: 4.603e+02
   cx: 3.630e+02
   cy: 2.481e+02
""",
    "kb": """%YAML:1.0
---
model_type: KANNALA_BRANDT
camera_name: camera
image_width: 512
image_height: 512
projection_parameters:
   k2: 0.0034823894022493434
   k3: 0.0007150348452162257
   k4: -0.0020532361418706202
   k5: 0.00020293673591811182
   mu: 190.97847715128717
   mv: 190.9733070521226
   u0: 254.93170605935475
   v0: 256.8974428996504
""",
    "mei": """%YAML:1.0
---
model_type: MEI
camera_name: cam0
image_width: 640
image_height: 480
mirror_parameters:
   xi: 1.8
distortion_parameters:
   k1: -0.2
   k2: 0.1
   p1: 1.0e-04
   p2: -2.0e-04
projection_parameters:
   gamma1: 700.5
   gamma2: 705.25
   u0: 320.5
   v0: 239.75
""",
    "mei_default": """%YAML:1.0
---
camera_name: no_model_type_means_mei
image_width: 640
image_height: 480
mirror_parameters:
   xi: 0.9
distortion_parameters:
   k1: 0.
   k2: 0.
   p1: 0.
   p2: 0.
projection_parameters:
   gamma1: 500.
   gamma2: 500.
   u0: 320.
   v0: 240.
""",
    "scaramuzza": """%YAML:1.0
---
model_type: scaramuzza
camera_name: ocam
image_width: 752
image_height: 480
poly_parameters:
   p0: -260.0
   p1: 0.0
   p2: 1.9e-03
   p3: -2.0e-06
   p4: 1.1e-08
inv_poly_parameters:
   p0: 380.0
   p1: 210.0
affine_parameters:
   ac: 1.0006
   ad: -2.0e-04
   ae: 3.0e-04
   cx: 376.2
   cy: 239.7
""",
}


def _product_vector(c):
    """CameraParams -> the parameter vector in the order Camera::writeParameters uses for that model"""
    from mobile_slam_b200 import _abi
    if c.model_type == _abi.VS_CAM_PINHOLE or c.model_type == _abi.VS_CAM_KANNALA_BRANDT:
        return list(c.d) + [c.fx, c.fy, c.cx, c.cy]
    if c.model_type == _abi.VS_CAM_MEI:
        return [c.xi] + list(c.d) + [c.fx, c.fy, c.cx, c.cy]
    return list(c.affine) + [c.cx, c.cy] + list(c.poly)        # C, D, E, center_x, center_y, poly[5] (+ inv_poly, unused)


@pytest.mark.parametrize("name", sorted(_YAMLS))
def test_read_intrinsic_parameter_matches_compiled_reference(ref_lib, tmp_path, name):
    from mobile_slam_b200.calib import camera_from_yaml
    p = tmp_path / f"{name}.yaml"
    p.write_text(_YAMLS[name])
    ref = RB.read_intrinsic_parameters(str(p))
    assert ref is not None
    mt, w, h, vec = ref
    c = camera_from_yaml(str(p))
    assert (c.model_type, c.width, c.height) == (mt, w, h)
    mine = _product_vector(c)
    assert np.array_equal(np.asarray(mine, np.float64), vec[:len(mine)]), (mine, vec)


def test_read_intrinsic_parameter_failures(ref_lib, tmp_path):
    from mobile_slam_b200.calib import camera_from_yaml
    bad = tmp_path / "bad.yaml"
    bad.write_text("%YAML:1.0\n---\nmodel_type: FISHEYE_X\nimage_width: 10\nimage_height: 10\n")
    assert RB.read_intrinsic_parameters(str(bad)) is None            # CameraFactory returns CameraPtr() (CameraFactory.cc:107-110)
    with pytest.raises(ValueError):
        camera_from_yaml(str(bad))
    assert RB.read_intrinsic_parameters(str(tmp_path / "missing.yaml")) is None


def test_reference_config_files(ref_lib):
    """the reference's own calibration files, where the reference tree is present"""
    import glob
    from mobile_slam_b200.calib import camera_from_yaml
    files = sorted(glob.glob("/root/reference/config/*.yaml"))
    if not files:
        pytest.skip("/root/reference/config absent")
    for f in files:
        ref = RB.read_intrinsic_parameters(f)
        assert ref is not None, f
        mt, w, h, vec = ref
        c = camera_from_yaml(f)
        mine = _product_vector(c)
        assert (c.model_type, c.width, c.height) == (mt, w, h) and np.array_equal(np.asarray(mine, np.float64), vec[:len(mine)]), f
