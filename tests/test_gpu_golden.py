"""GPU parity against the committed golden vectors (tests/golden/*.npz, OpenCV 4.13.0 outputs): needs neither cv2
nor /root/reference at run time.  Bars: CLAHE / pyramid bit-exact; corners identical; LK status equal and positions
within 0.01 px; sequence ids / track_cnt identical and positions within 0.01 px (north_star tolerances)."""
import importlib.util
import os

import numpy as np
import pytest

from mobile_slam_b200 import BatchTracker, FeatureTracker, TrackerConfig, CameraParams, VS_RANSAC_OFF, config_from_synth
from mobile_slam_b200 import _abi

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _mk():
    mk = importlib.util.spec_from_file_location("make_golden_cfg", os.path.join(GOLD, "make_golden.py"))
    src = open(os.path.join(GOLD, "make_golden.py")).read()
    ns = {}
    # only the two config dicts are needed (the module itself imports cv2)
    start = src.index("SEQ_CFG = ")
    end = src.index("def stages")
    from mobile_slam_b200.synth import CONFIGS
    exec(src[start:end], {"CONFIGS": CONFIGS}, ns)
    return ns


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLD, "stages_small.npz"))


@pytest.fixture(scope="module")
def small(built_lib, g):
    h, w = g["frame0"].shape
    tc = TrackerConfig(col=w, row=h, max_cnt=40, min_dist=10, ransac_mode=VS_RANSAC_OFF, capacity=256)
    t = BatchTracker(tc, CameraParams(width=w, height=h, cx=w / 2, cy=h / 2), 1)
    yield t
    t.close()


def test_clahe_golden(small, g):
    assert np.array_equal(small.stage_clahe(g["frame0"]), g["clahe0"])
    assert np.array_equal(small.stage_clahe(g["frame1"]), g["clahe1"])


def test_clahe_padding_golden(built_lib):
    # cv::CLAHE's padding rule (a side on the 8 x 8 tile grid still grows by 8 when the other side is off it)
    c = np.load(os.path.join(GOLD, "clahe_padding.npz"))
    for k in [k for k in c.files if k.startswith("in_")]:
        img, ref = c[k], c["out_" + k[3:]]
        h, w = img.shape
        tc = TrackerConfig(col=w, row=h, max_cnt=40, min_dist=10, ransac_mode=VS_RANSAC_OFF, capacity=256)
        t = BatchTracker(tc, CameraParams(width=w, height=h, cx=w / 2, cy=h / 2), 1)
        got = t.stage_clahe(img)
        t.close()
        assert np.array_equal(got, ref), (k, int(np.count_nonzero(got != ref)))


def test_pyramid_golden(small, g):
    lv = small.stage_pyramid(g["clahe0"])
    assert len(lv) == int(g["n_levels"])
    for l, a in enumerate(lv):
        assert np.array_equal(a, g[f"pyr{l}"]), l


def test_lk_golden(small, g):
    got, st = small.stage_lk(g["clahe0"], g["clahe1"], g["lk_pts"])
    assert np.array_equal(st, g["lk_status"])
    ok = st == 1
    assert np.abs(got[ok] - g["lk_next"][ok]).max() <= 0.01


def test_good_features_golden(small, g):
    got = small.stage_good_features(g["clahe0"], np.zeros((0, 2), np.float32), 40)
    assert np.array_equal(got, g["gftt_all"])
    got, mask = small.stage_good_features(g["clahe0"], g["gftt_kept"], 25, return_mask=True)
    assert np.array_equal(mask, g["gftt_mask"])
    assert np.array_equal(got, g["gftt_masked"])


def test_min_eig_golden(small, g):
    e = small.stage_min_eig(g["clahe0"])
    ref = g["min_eig"]
    assert np.percentile(np.abs(e - ref) / np.maximum(np.abs(ref), 1e-12), 99) < 1e-5


@pytest.mark.parametrize("fname,key", [("sequence_small.npz", "SEQ_CFG"), ("sequence_fisheye.npz", "SEQ_FISH_CFG")])
def test_sequence_golden(built_lib, fname, key):
    q = np.load(os.path.join(GOLD, fname))
    cfg = _mk()[key]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    ft = FeatureTracker(tc)
    ft.setCamera(cam)
    n_ref = n_common = 0
    for k in range(int(q["n_frames"])):
        ft.detectAndTrack(q[f"frame{k}"], 0.05 * k)
        i = 0
        while ft.updateID(i):
            i += 1
        rid = q[f"ids{k}"].tolist()
        gi = {v: j for j, v in enumerate(ft.ids)}
        common = [v for v in rid if v in gi]
        n_ref += len(rid)
        n_common += len(common)
        a = np.array([gi[v] for v in common])
        b = np.array([rid.index(v) for v in common])
        assert np.abs(np.asarray(ft.cur_pts)[a] - q[f"pts{k}"][b]).max() <= 0.01
        assert np.abs(np.asarray(ft.cur_undistorted_pts)[a] - q[f"un{k}"][b]).max() <= 1e-4
        assert np.array_equal(np.asarray(ft.track_cnt)[a], q[f"cnt{k}"][b])
        dv = np.abs(np.asarray(ft.pts_velocity)[a] - q[f"vel{k}"][b]).max()
        assert dv <= 2 * 1e-4 / 0.05, dv          # velocity = delta undist / dt: what the 1e-4 undistortion bound implies
    assert n_common >= 0.99 * n_ref, (n_common, n_ref)


def test_two_pixel_lane_variant_golden(built_lib):
    """shi_tomasi_kernel<2, .> (VS_ST_PX=2, read once per process): the same golden corner lists, eig map and
    free-running sequences through the two-pixels-per-lane build of the strip walk."""
    import subprocess
    import sys
    if os.environ.get("VS_ST_PX") == "2":
        pytest.skip("already inside the VS_ST_PX=2 run")
    env = dict(os.environ, VS_ST_PX="2")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-m", "gpu", "-q", "-x", "-k",
                        "good_features_golden or min_eig_golden or sequence_golden"],
                       env=env, capture_output=True, text=True, timeout=600, cwd=os.path.dirname(GOLD))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert " passed" in r.stdout
