"""Edge cases of the path through the C ABI, against the reference's OpenCV CPU tracker on identical input:
row-strided / unaligned caller buffers, featureless frames, scene cuts (every track lost at once), repeated
timestamps, and a frame that is smaller than one LK window at the top pyramid level.
"""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from mobile_slam_b200 import BatchTracker, VS_RANSAC_OFF, config_from_synth  # noqa: E402
from mobile_slam_b200.synth import CONFIGS, SynthStream  # noqa: E402
from oracle import camera as ocam  # noqa: E402
from oracle import cv_reference as R  # noqa: E402

pytestmark = pytest.mark.gpu

SMALL = dict(width=333, height=251, max_cnt=100, min_dist=15, max_level=3, fisheye=0, model="PINHOLE",
             intr=(300.0, 300.0, 166.0, 125.0), dist=(-0.1, 0.01, 1e-4, -1e-4), focal_length=460.0)


def _run_both(cfg, frames, times):
    """ids / track_cnt / points / velocities per frame from the CUDA tracker and from the reference restatement."""
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    bt = BatchTracker(tc, cam, 1)
    ref = R.RefFeatureTracker(R.config_from_synth(cfg, use_ransac=False), ocam.from_config(cfg))
    out = []
    for f, t in zip(frames, times):
        r = bt.track([f], [t])[0]
        ref.detect_and_track(np.ascontiguousarray(f), t)
        ref.update_all_ids()
        out.append((r, list(ref.ids), list(ref.track_cnt), np.array(ref.cur_pts, np.float32).reshape(-1, 2),
                    np.array(ref.pts_velocity, np.float32).reshape(-1, 2)))
    bt.close()
    return out


def _assert_same(out, pos_tol=0.01, min_exact=None):
    exact = 0
    for k, (r, ids, cnt, pts, vel) in enumerate(out):
        same = r.ids.tolist() == ids and r.track_cnt.tolist() == cnt
        exact += same
        if same and len(ids):
            assert np.abs(r.cur_pts - pts).max() <= pos_tol, k
    need = len(out) if min_exact is None else min_exact
    assert exact >= need, f"{exact}/{len(out)} frames identical in ids + track_cnt"


def test_row_strided_and_unaligned_frames_match_contiguous(built_lib):
    cfg = SMALL
    W, H = cfg["width"], cfg["height"]
    st = SynthStream(5, W, H)
    frames = [st.frame(k) for k in range(5)]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    a, b = BatchTracker(tc, cam, 1), BatchTracker(tc, cam, 1)
    for k, f in enumerate(frames):
        wide = np.zeros((H, W + 29), np.uint8)            # step 362: rows neither 4- nor 16-byte aligned
        wide[:, 3:3 + W] = f                              # first pixel at an odd address
        view = wide[:, 3:3 + W]
        assert not view.flags["C_CONTIGUOUS"] and view.strides == (W + 29, 1)
        ra = a.track([f], [0.05 * k])[0]
        rb = b.track([view], [0.05 * k])[0]
        assert ra.ids.tolist() == rb.ids.tolist() and ra.track_cnt.tolist() == rb.track_cnt.tolist()
        assert np.array_equal(ra.cur_pts, rb.cur_pts) and np.array_equal(ra.pts_velocity, rb.pts_velocity)
    a.close()
    b.close()


def test_featureless_frames_then_texture(built_lib):
    cfg = CONFIGS["C2"]
    W, H = cfg["width"], cfg["height"]
    st = SynthStream(9, W, H)
    flat = np.full((H, W), 117, np.uint8)
    frames = [flat, flat, st.frame(0), st.frame(1), flat, st.frame(2), st.frame(3)]
    out = _run_both(cfg, frames, [0.05 * k for k in range(len(frames))])
    assert len(out[0][0].ids) == 0 and len(out[1][0].ids) == 0          # nothing to detect on a constant image
    assert len(out[3][0].ids) > 100
    _assert_same(out)


def test_scene_cut_loses_every_track(built_lib):
    cfg = CONFIGS["C1"]
    W, H = cfg["width"], cfg["height"]
    a, b = SynthStream(40, W, H), SynthStream(41, W, H)
    frames = [a.frame(k) for k in range(4)] + [b.frame(k) for k in range(4)]
    out = _run_both(cfg, frames, [0.05 * k for k in range(len(frames))])
    # LK on unrelated images is arbitrary but deterministic arithmetic: the two implementations may disagree on a
    # borderline status flag right at the cut, never on the frames before it or on the recovery after it
    _assert_same(out[:4])
    n_before = max(out[3][0].ids.tolist()) + 1
    fresh_at_cut = sum(1 for i in out[4][0].ids.tolist() if i >= n_before)
    assert fresh_at_cut >= 30, fresh_at_cut                               # most tracks died at the cut and were replaced
    _assert_same(out, min_exact=6)


def test_repeated_and_decreasing_timestamps_give_zero_velocity(built_lib):
    cfg = CONFIGS["C2"]
    st = SynthStream(12, cfg["width"], cfg["height"])
    frames = [st.frame(k) for k in range(6)]
    times = [0.0, 0.05, 0.10, 0.10, 0.10 + 5e-5, 0.08]                  # dt = 0, dt < 1e-4, dt < 0 (feature_tracker.cpp:360)
    out = _run_both(cfg, frames, times)
    _assert_same(out)
    for k in (3, 4, 5):
        r, _, _, _, vel = out[k]
        assert np.all(r.pts_velocity == 0.0) and np.all(vel == 0.0), k
    r, _, cnt, _, vel = out[2]
    assert np.any(r.pts_velocity != 0.0)
    old = np.array(cnt) >= 3
    assert np.abs(r.pts_velocity[old] - vel[old]).max() <= 0.01 / 0.05 * 2


def test_tiny_frame_fewer_pyramid_levels(built_lib):
    # 120 x 90 with a 21 px window and maxLevel 3: the pyramid stops at level 2 (30 x 23; the next, 15 x 12, is not
    # larger than the window, S2), and CLAHE equalises it as 128 x 96 (width already a multiple of 8, height not)
    cfg = dict(SMALL, width=120, height=90, max_cnt=40, min_dist=8, intr=(100.0, 100.0, 60.0, 45.0))
    st = SynthStream(3, 200, 200)
    frames = [np.ascontiguousarray(st.frame(k)[40:130, 30:150]) for k in range(5)]
    out = _run_both(cfg, frames, [0.05 * k for k in range(5)])
    assert len(out[0][0].ids) > 5
    _assert_same(out)
