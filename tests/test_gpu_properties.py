"""Size-independent properties of the path at BASELINE.json's full sizes (no oracle needed, so they run at scale):
C5 (64 independent 512x512 streams) and C4 (1280x720, 2000 features).  What must hold for ANY input:
  * a stream's tracks do not depend on which batch / lane / handle it runs in (no cross-stream state);
  * ids are unique per stream, new ids are assigned in increasing order, track_cnt of a surviving id grows by exactly 1;
  * every kept point lies inside the 1-px border (inBorder, feature_tracker.cpp:12-18) and inside the fisheye circle;
  * after setMask + goodFeaturesToTrack no kept point lies inside an earlier-ranked kept point's cv::circle(min_dist) disc
    (feature_tracker.cpp:79-86), and newly detected corners are >= min_dist apart (goodFeaturesToTrack's grid rule);
  * velocity is exactly 0 for features with track_cnt < 3 (the id-map quirk, feature_tracker.cpp:351-378);
  * re-running the same frames gives bit-identical results (determinism, incl. the persistent-warp LK work queue).
"""
import numpy as np
import pytest

from mobile_slam_b200 import BatchTracker, VS_RANSAC_DEVICE, VS_RANSAC_OFF, config_from_synth
from mobile_slam_b200.synth import CONFIGS, SynthStream

pytestmark = pytest.mark.gpu


def _half_widths(r):
    hw = np.full(r + 1, -1, np.int64)
    dx, dy, err, plus, minus = r, 0, 0, 1, (r << 1) - 1
    while dx >= dy:
        hw[dy] = max(hw[dy], dx)
        hw[dx] = max(hw[dx], dy)
        dy += 1
        err += plus
        plus += 2
        m = -1 if err > 0 else 0
        err -= minus & m
        dx += m
        minus -= m & 2
    return hw


def _check_frame(cfg, prev, cur, k):
    W, H, R = cfg["width"], cfg["height"], cfg["min_dist"]
    ids, cnt, pts = cur.ids, cur.track_cnt, cur.cur_pts
    assert len(np.unique(ids)) == len(ids)
    assert len(ids) <= cfg["max_cnt"]
    x, y = np.rint(pts[:, 0]).astype(np.int64), np.rint(pts[:, 1]).astype(np.int64)
    assert np.all((x >= 1) & (x < W - 1) & (y >= 1) & (y < H - 1))
    if cfg["fisheye"]:
        cx, cy, r = int(W / 2.0), int(H / 2.0), int(0.95 * min(W / 2.0, H / 2.0))
        fh = _half_widths(r)
        dy = np.abs(y - cy)
        assert np.all(dy <= r) and np.all(np.abs(x - cx) <= fh[np.minimum(dy, r)])
    n_old = len(ids) - cur.n_new
    # old points come first, ordered by track_cnt descending (setMask's sort), new ones after with count 1 and fresh ids
    assert np.all(np.diff(cnt[:n_old]) <= 0) and np.all(cnt[n_old:] == 1)
    if cur.n_new > 1:
        assert np.all(np.diff(ids[n_old:]) == 1)
    if prev is not None:
        pc = dict(zip(prev.ids.tolist(), prev.track_cnt.tolist()))
        for i, c in zip(ids[:n_old].tolist(), cnt[:n_old].tolist()):
            assert pc[i] + 1 == c
        if cur.n_new:
            assert ids[n_old] > prev.ids.max()
    # disc exclusion among tracked points (rank order = array order) and against new corners
    hw = _half_widths(R)
    dxm = np.abs(x[:, None] - x[None, :])
    dym = np.abs(y[:, None] - y[None, :])
    inside = (dym <= R) & (dxm <= hw[np.minimum(dym, R)])
    np.fill_diagonal(inside, False)
    earlier = np.tril(np.ones_like(inside), -1).astype(bool)          # [q, a] with a < q
    old_rows = np.arange(len(ids)) < n_old
    assert not np.any(inside & earlier & old_rows[None, :]), "a point lies inside an earlier kept point's disc"
    if cur.n_new > 1:
        d2 = (dxm ** 2 + dym ** 2)[n_old:, n_old:]
        np.fill_diagonal(d2, 10 ** 9)
        assert d2.min() >= R * R
    assert np.all(cur.pts_velocity[cnt < 3] == 0)
    assert np.all(np.isfinite(cur.cur_undistorted_pts)) and np.all(np.isfinite(cur.pts_velocity))


@pytest.mark.parametrize("mode", [VS_RANSAC_OFF, VS_RANSAC_DEVICE], ids=["ransac_off", "ransac_device"])
def test_c5_64_streams_invariants_and_independence(built_lib, mode):
    # with the device F-matrix RANSAC too: its sample sequence is seeded from the stream's own points, so a stream's
    # result does not depend on where it sits in a batch, a lane or a handle
    cfg = CONFIGS["C1"]
    S, F = 64, 8
    streams = [SynthStream(700 + s, cfg["width"], cfg["height"], motion=1.0 + (s % 4)) for s in range(S)]
    frames = [np.stack([st.frame(k) for st in streams]) for k in range(F)]
    tc, cam = config_from_synth(cfg, ransac_mode=mode)
    tc8, _ = config_from_synth(cfg, ransac_mode=mode, pipeline_lanes=8)
    whole, lanes = BatchTracker(tc, cam, S), BatchTracker(tc8, cam, S)
    halves = [BatchTracker(tc, cam, S // 2), BatchTracker(tc, cam, S // 2)]     # "2 GPUs": streams s mod 2
    prev = [None] * S
    for k in range(F):
        ts = [0.05 * k] * S
        a = whole.track(list(frames[k]), ts)
        b = lanes.track(list(frames[k]), ts)
        h = [halves[g].track([frames[k][s] for s in range(g, S, 2)], ts[:S // 2]) for g in range(2)]
        for s in range(S):
            c = h[s % 2][s // 2]
            for other in (b[s], c):
                assert np.array_equal(a[s].ids, other.ids) and np.array_equal(a[s].track_cnt, other.track_cnt)
                assert np.array_equal(a[s].cur_pts, other.cur_pts) and np.array_equal(a[s].pts_velocity, other.pts_velocity)
            _check_frame(cfg, prev[s], a[s], k)
            prev[s] = a[s]
    assert np.mean([np.sum(p.track_cnt > 1) for p in prev]) > 100      # the path really tracked
    for t in [whole, lanes] + halves:
        t.close()


@pytest.mark.parametrize("mode", [VS_RANSAC_OFF, VS_RANSAC_DEVICE], ids=["ransac_off", "ransac_device"])
def test_c4_2000_features_invariants_and_determinism(built_lib, mode):
    cfg = CONFIGS["C4"]
    st = SynthStream(900, cfg["width"], cfg["height"], motion=2.0)
    frames = [st.frame(k) for k in range(5)]
    tc, cam = config_from_synth(cfg, ransac_mode=mode)
    runs = []
    for rep in range(2):
        bt = BatchTracker(tc, cam, 1)
        out, prev = [], None
        for k, f in enumerate(frames):
            r = bt.track([f], [0.05 * k])[0]
            if rep == 0:
                _check_frame(cfg, prev, r, k)
            prev = r
            out.append(r)
        runs.append(out)
        bt.close()
    for a, b in zip(*runs):
        assert np.array_equal(a.ids, b.ids) and np.array_equal(a.cur_pts, b.cur_pts) and np.array_equal(a.track_cnt, b.track_cnt)
    assert len(runs[0][-1].ids) > 1500 and np.sum(runs[0][-1].track_cnt > 1) > 1200


def test_two_gpus_in_one_process(built_lib):
    # vs::ShardedTracker's layout: one handle per GPU inside ONE process.  Per-device kernel attributes (dynamic shared
    # memory opt-in for the large-capacity glue kernels) must be set on every device, and a stream must give the same
    # tracks whichever GPU it runs on.
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cfg = dict(CONFIGS["C4"])
    st = SynthStream(901, cfg["width"], cfg["height"], motion=2.0)
    frames = [st.frame(k) for k in range(4)]
    out = []
    for dev in (0, 1):
        tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_DEVICE, device=dev)
        out.append((BatchTracker(tc, cam, 1), []))
    for k, f in enumerate(frames):                      # interleaved: both devices live at the same time
        for bt, res in out:
            res.append(bt.track([f], [0.05 * k])[0])
    for a, b in zip(out[0][1], out[1][1]):
        assert np.array_equal(a.ids, b.ids) and np.array_equal(a.track_cnt, b.track_cnt)
        assert np.array_equal(a.cur_pts, b.cur_pts) and np.array_equal(a.pts_velocity, b.pts_velocity)
    assert len(out[0][1][-1].ids) > 1500
    for bt, _ in out:
        bt.close()
