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
    assert np.abs(r.pts_velocity[old] - vel[old]).max() <= 2 * 1e-4 / 0.05   # 1e-4 undistortion bound, twice, / dt


def test_tiny_frame_fewer_pyramid_levels(built_lib):
    # 120 x 90 with a 21 px window and maxLevel 3: the pyramid stops at level 2 (30 x 23; the next, 15 x 12, is not
    # larger than the window, S2), and CLAHE equalises it as 128 x 96 (width already a multiple of 8, height not)
    cfg = dict(SMALL, width=120, height=90, max_cnt=40, min_dist=8, intr=(100.0, 100.0, 60.0, 45.0))
    st = SynthStream(3, 200, 200)
    frames = [np.ascontiguousarray(st.frame(k)[40:130, 30:150]) for k in range(5)]
    out = _run_both(cfg, frames, [0.05 * k for k in range(5)])
    assert len(out[0][0].ids) > 5
    _assert_same(out)


def test_runtime_mutators_match_reference_config_changes(built_lib):
    # VIOEngine::setMobileParams / setTrackingParams rewrite g_config between frames (vio_engine.cpp:354-389); the
    # handle's vs_tracker_set_params must have the same effect on the very next frame
    cfg = CONFIGS["C2"]
    st = SynthStream(23, cfg["width"], cfg["height"], motion=1.5)
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF, capacity=512)
    bt = BatchTracker(tc, cam, 1)
    ref = R.RefFeatureTracker(R.config_from_synth(cfg, use_ransac=False), ocam.from_config(cfg))
    changes = {3: dict(max_cnt=200, min_dist=20), 5: dict(lk_window=15, lk_pyramid=2, lk_iterations=20, lk_eps=0.03),
               7: dict(max_cnt=100, min_dist=35), 9: dict(lk_window=31, lk_pyramid=4)}
    names = dict(max_cnt="max_cnt", min_dist="min_dist", lk_window="lk_window_size", lk_pyramid="lk_pyramid_levels",
                 lk_iterations="lk_iterations", lk_eps="lk_eps")
    exact = 0
    for k in range(12):
        if k in changes:
            bt.set_params(**changes[k])
            for a, v in changes[k].items():
                setattr(ref.cfg, names[a], v)
        f = st.frame(k)
        r = bt.track([f], [0.05 * k])[0]
        ref.detect_and_track(f, 0.05 * k)
        ref.update_all_ids()
        same = r.ids.tolist() == list(ref.ids) and r.track_cnt.tolist() == list(ref.track_cnt)
        exact += same
        if same:
            assert np.abs(r.cur_pts - np.array(ref.cur_pts, np.float32).reshape(-1, 2)).max() <= 0.01, k
        if k == 4:
            assert 150 < len(r.ids) <= 200
    assert exact >= 11, f"{exact}/12 frames identical"
    bt.close()


def test_callback_ransac_on_a_multi_lane_batch(built_lib):
    # the host hop of VS_RANSAC_CALLBACK with several streams split over pipeline lanes: each stream's callback sees
    # its own points, and the batch equals the same streams tracked one handle each
    from mobile_slam_b200 import VS_RANSAC_CALLBACK
    cfg = dict(CONFIGS["C2"])
    S = 4
    streams = [SynthStream(500 + s, cfg["width"], cfg["height"], motion=2.0) for s in range(S)]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_CALLBACK)
    tcl, _ = config_from_synth(cfg, ransac_mode=VS_RANSAC_CALLBACK, pipeline_lanes=2)
    seen = {"batch": set(), "single": 0}

    def make(tag):
        def reject(stream, a, b, px):
            if tag == "batch":
                seen["batch"].add(stream)
            else:
                seen["single"] += 1
            _, status = cv2.findFundamentalMat(a, b, cv2.FM_RANSAC, 1.0, 0.99)
            return np.ones(len(a), np.uint8) if status is None else status.reshape(-1)
        return reject
    batch = BatchTracker(tcl, cam, S)
    batch.set_reject_callback(make("batch"))
    singles = [BatchTracker(tc, cam, 1) for _ in range(S)]
    for t in singles:
        t.set_reject_callback(make("single"))
    for k in range(6):
        frames = [st.frame(k) for st in streams]
        rb = batch.track(frames, [0.05 * k] * S)
        for s in range(S):
            r1 = singles[s].track([frames[s]], [0.05 * k])[0]
            assert rb[s].ids.tolist() == r1.ids.tolist() and rb[s].track_cnt.tolist() == r1.track_cnt.tolist(), (k, s)
            assert np.array_equal(rb[s].cur_pts, r1.cur_pts)
    assert seen["batch"] == set(range(S)) and seen["single"] > 0
    batch.close()
    for t in singles:
        t.close()


def test_device_ransac_and_fb_check_through_submit_wait_and_lanes(built_lib):
    # the asynchronous pair and the lane split must not change what the device RANSAC (per-frame seed) and the
    # forward/backward check produce
    from mobile_slam_b200 import VS_RANSAC_DEVICE
    cfg = CONFIGS["C1"]
    S = 4
    streams = [SynthStream(700 + s, cfg["width"], cfg["height"], motion=2.0) for s in range(S)]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_DEVICE, fb_check=1)
    tcl, _ = config_from_synth(cfg, ransac_mode=VS_RANSAC_DEVICE, fb_check=1, pipeline_lanes=2)
    sync, asyn = BatchTracker(tc, cam, S), BatchTracker(tcl, cam, S)
    frames = [[st.frame(k) for st in streams] for k in range(8)]
    got = []
    asyn.submit(frames[0], [0.0] * S)
    for k in range(1, 8):
        asyn.submit(frames[k], [0.05 * k] * S)
        got.append(asyn.wait())
    got.append(asyn.wait())
    ran = 0
    for k in range(8):
        rs = sync.track(frames[k], [0.05 * k] * S)
        for s in range(S):
            assert rs[s].ids.tolist() == got[k][s].ids.tolist() and rs[s].track_cnt.tolist() == got[k][s].track_cnt.tolist(), (k, s)
            assert np.array_equal(rs[s].cur_pts, got[k][s].cur_pts) and np.array_equal(rs[s].pts_velocity, got[k][s].pts_velocity)
            ran += rs[s].counters["ransac_ran"]
    assert ran >= 12
    sync.close()
    asyn.close()


def test_fisheye_mask_image_with_grey_levels(built_lib):
    # feature_tracker.cpp:42-46: a mask image replaces the generated circle.  setMask keeps a tracked point only where
    # the mask is exactly 255 (:80), goodFeaturesToTrack detects wherever it is non-zero: a grey (1..254) region
    # therefore gets new corners every frame and loses them again on the next one
    cfg = dict(CONFIGS["C1"])
    W, H = cfg["width"], cfg["height"]
    m = np.full((H, W), 255, np.uint8)
    m[:, :80] = 0                       # nothing on the left
    m[:120, :] = 100                    # grey band on top
    m[300:360, 200:420] = 0             # a hole
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF, fisheye_mask=m)
    bt = BatchTracker(tc, cam, 1)
    ref = R.RefFeatureTracker(R.config_from_synth(cfg, use_ransac=False), ocam.from_config(cfg))
    ref.fisheye_mask = m.copy()
    st = SynthStream(61, W, H)
    exact, grey_new = 0, 0
    for k in range(8):
        f = st.frame(k)
        r = bt.track([f], [0.05 * k])[0]
        ref.detect_and_track(f, 0.05 * k)
        ref.update_all_ids()
        same = r.ids.tolist() == list(ref.ids) and r.track_cnt.tolist() == list(ref.track_cnt)
        exact += same
        if same and len(r.ids):
            assert np.abs(r.cur_pts - np.array(ref.cur_pts, np.float32).reshape(-1, 2)).max() <= 0.01, k
        x, y = np.rint(r.cur_pts[:, 0]).astype(int), np.rint(r.cur_pts[:, 1]).astype(int)
        assert np.all(m[y, x] != 0), k
        old = r.track_cnt > 1
        assert np.all(m[y[old], x[old]] == 255), k                      # survivors sit on 255 only
        grey_new += int(np.sum(m[y[~old], x[~old]] == 100))
    assert exact >= 7, exact
    assert grey_new > 0
    bt.close()


def test_device_ransac_edge_recovery(built_lib):
    # feature_tracker.cpp:236-285: with f_threshold_edge_factor > 0 a point rejected by the RANSAC mask is restored when
    # it lies in the outer image region (r_ratio >= 0.3) and within F_threshold * (1 + factor * r_ratio^2) of its
    # epipolar line.  Same stream through two handles (factor 0 and 3): while both still hold the same features the
    # second one keeps a superset, and what it restores lies in the outer region.
    from mobile_slam_b200 import VS_RANSAC_DEVICE
    cfg = CONFIGS["C1"]
    W, H = cfg["width"], cfg["height"]
    st = SynthStream(77, W, H, motion=2.0)
    tc0, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_DEVICE)
    tc3, _ = config_from_synth(cfg, ransac_mode=VS_RANSAC_DEVICE, f_threshold_edge_factor=3.0)
    a, b = BatchTracker(tc0, cam, 1), BatchTracker(tc3, cam, 1)
    prev_same, restored, compared = True, 0, 0
    r_max = np.hypot(W / 2.0, H / 2.0)
    for k in range(40):
        f = st.frame(k)
        ra, rb = a.track([f], [0.05 * k])[0], b.track([f], [0.05 * k])[0]
        if prev_same and ra.counters["ransac_ran"]:
            assert rb.counters["ransac_ran"]
            compared += 1
            assert rb.counters["n_ransac_in"] == ra.counters["n_ransac_in"]
            assert rb.counters["n_ransac_out"] >= ra.counters["n_ransac_out"], k
            old_a = set(ra.ids[ra.track_cnt > 1].tolist())
            extra = [i for i, c in zip(rb.ids.tolist(), rb.track_cnt.tolist()) if c > 1 and i not in old_a]
            for i in extra:
                p = rb.cur_pts[rb.ids.tolist().index(i)]
                assert np.hypot(p[0] - W / 2.0, p[1] - H / 2.0) / r_max >= 0.3 - 1e-6, (k, i, p)
            restored += rb.counters["n_ransac_out"] - ra.counters["n_ransac_out"]
        prev_same = prev_same and ra.ids.tolist() == rb.ids.tolist()
        if not prev_same:
            break
    assert compared >= 1
    assert restored >= 1, "edge recovery never restored a point on this sequence"
    a.close()
    b.close()


def test_ragged_batch_blank_and_busy_streams_together(built_lib):
    # one launch covers streams in very different states: one that is featureless for 5 frames and then gets texture,
    # one with a scene cut, a normal one, one that never has a feature.  Each must equal the same stream tracked alone.
    cfg = dict(CONFIGS["C2"])
    W, H = cfg["width"], cfg["height"]
    a, b, c, d = (SynthStream(s, W, H, motion=2.0) for s in (810, 811, 812, 813))
    flat = np.full((H, W), 90, np.uint8)

    def frames(k):
        return [flat if k < 5 else a.frame(k), (b.frame(k) if k < 4 else c.frame(k)), d.frame(k), flat]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    batch = BatchTracker(tc, cam, 4)
    singles = [BatchTracker(tc, cam, 1) for _ in range(4)]
    for k in range(8):
        fr = frames(k)
        rb = batch.track(fr, [0.05 * k] * 4)
        for s in range(4):
            r1 = singles[s].track([fr[s]], [0.05 * k])[0]
            assert rb[s].ids.tolist() == r1.ids.tolist() and rb[s].track_cnt.tolist() == r1.track_cnt.tolist(), (k, s)
            assert np.array_equal(rb[s].cur_pts, r1.cur_pts) and np.array_equal(rb[s].pts_velocity, r1.pts_velocity)
        assert len(rb[3].ids) == 0
        assert (len(rb[0].ids) == 0) == (k < 5)
    assert len(rb[1].ids) > 100 and len(rb[2].ids) > 100
    batch.close()
    for t in singles:
        t.close()


def test_unsynchronised_device_calls_keep_their_timestamps(built_lib):
    """vs_tracker_track_batch_device(results_to_host = 0) returns without a sync; back-to-back calls must still see
    their OWN timestamps (stream-ordered timestamp ring), i.e. velocities equal the synchronous path's."""
    torch = pytest.importorskip("torch")
    cfg = dict(CONFIGS["C2"], width=376, height=240, max_cnt=100, min_dist=15, intr=(230.0, 230.0, 188.0, 120.0))
    n_streams, n_frames = 3, 14               # more frames than ring slots
    W, H = cfg["width"], cfg["height"]
    streams = [SynthStream(60 + s, W, H, motion=1.5) for s in range(n_streams)]
    frames = np.stack([np.stack([s.frame(k) for s in streams]) for k in range(n_frames)])     # (k, s, H, W)
    times = [[0.05 * k + 0.011 * s * k for s in range(n_streams)] for k in range(n_frames)]   # uneven dt per stream
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    d = torch.from_numpy(frames).cuda()
    sync = BatchTracker(tc, cam, n_streams)
    want = [sync.track_device(d[k].data_ptr(), W, H * W, times[k], to_host=True) for k in range(n_frames)]
    sync.close()
    fast = BatchTracker(tc, cam, n_streams)
    for k in range(n_frames):
        fast.track_device(d[k].data_ptr(), W, H * W, times[k], to_host=False)
    got = fast.fetch()
    for s in range(n_streams):
        assert np.array_equal(got[s].ids, want[-1][s].ids)
        assert np.array_equal(got[s].cur_pts, want[-1][s].cur_pts)
        assert np.array_equal(got[s].pts_velocity, want[-1][s].pts_velocity)
        assert np.any(got[s].pts_velocity != 0.0)
    fast.close()
