"""Out-of-bounds-write detector of the repo's own: every device buffer of a handle created with VS_DEBUG_GUARD=1 is
bracketed by two 4 KiB bands of 0xA5 (tracker_state.h, DeviceArena); after the frame path, the stage entry points, ragged
batches, pipeline lanes and odd image sizes have run, no guard byte may have changed — and the results must equal those of
an unguarded handle on the same input (a guard only moves buffers, never changes arithmetic).

compute-sanitizer is not available on the GPU pool, so this is the check that the kernels' stores stay inside their
buffers (loads past the end fall into the 256-byte slack / the guard and are harmless by construction)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _sequence(name, n_streams, n_frames, mode, lanes=0, fb=0, sizes=None):
    from mobile_slam_b200 import BatchTracker, config_from_synth
    from mobile_slam_b200.synth import CONFIGS, SynthStream
    cfg = dict(CONFIGS[name])
    if sizes:
        cfg.update(width=sizes[0], height=sizes[1], intr=(0.8 * sizes[0], 0.8 * sizes[0], sizes[0] / 2, sizes[1] / 2))
    tc, cam = config_from_synth(cfg, ransac_mode=mode, pipeline_lanes=lanes, fb_check=fb)
    t = BatchTracker(tc, cam, n_streams)
    W, H = cfg["width"], cfg["height"]
    streams = [SynthStream(100 + s, max(W, 256), max(H, 256), motion=1.0 + 0.5 * (s % 3)) for s in range(n_streams)]
    out = []
    for k in range(n_frames):
        frames = [np.ascontiguousarray(s.frame(k)[:H, :W]) for s in streams]      # small sizes: a crop of a 256-px scene
        if n_streams > 2:                      # ragged: one blank stream, one scene cut
            frames[1] = np.zeros_like(frames[1])
            if k == 2:
                frames[2] = np.ascontiguousarray(streams[0].frame(40)[:H, :W])
        res = t.track(frames, [0.05 * k] * n_streams)
        out.append([(r.ids.copy(), r.cur_pts.copy(), r.track_cnt.copy(), r.cur_undistorted_pts.copy(), r.pts_velocity.copy()) for r in res])
    g = t.debug_guard_check()
    t.close()
    return out, g


def _stages(name):
    import cv2
    from mobile_slam_b200 import BatchTracker, VS_RANSAC_OFF, config_from_synth
    from mobile_slam_b200.synth import CONFIGS, SynthStream
    cfg = CONFIGS[name]
    w, h = cfg["width"], cfg["height"]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF, capacity=cfg["max_cnt"] + 64)
    t = BatchTracker(tc, cam, 1)
    s = SynthStream(7, w, h)
    a, b = s.frame(3), s.frame(4)
    rng = np.random.default_rng(3)
    pts = cv2.goodFeaturesToTrack(a, cfg["max_cnt"], 0.01, cfg["min_dist"]).reshape(-1, 2)
    edge = np.array([[0, 0], [w - 1, h - 1], [0.4, h - 1.2], [w - 1.3, 0.6], [w / 2, 0], [0, h / 2]], np.float32)
    pts = np.concatenate([pts, edge])[: t.capacity]
    out = [t.stage_clahe(a)]
    out += list(t.stage_pyramid(a))
    out += list(t.stage_lk(a, b, pts))
    out.append(t.stage_min_eig(a))
    out.append(t.stage_good_features(a, pts[:40], cfg["max_cnt"]))
    c2, m2 = t.stage_good_features(a, pts[:40], cfg["max_cnt"], return_mask=True)
    out += [c2, m2]
    flat = np.full((h, w), 90, np.uint8)                               # plateau: every pixel a candidate
    out.append(t.stage_good_features(flat, np.zeros((0, 2), np.float32), cfg["max_cnt"]))
    out.append(t.stage_undistort(pts))
    out.append(t.stage_set_mask(pts, rng.integers(1, 30, len(pts)).astype(np.int32)))
    nxt = pts + rng.normal(0, 0.5, pts.shape).astype(np.float32)
    out.append(t.stage_ransac(pts, nxt))
    g = t.debug_guard_check()
    t.close()
    return out, g


def _child():
    from mobile_slam_b200 import VS_RANSAC_DEVICE, VS_RANSAC_OFF
    import pickle
    res = {}
    res["seq_c1"] = _sequence("C1", 1, 5, VS_RANSAC_DEVICE)
    res["seq_c2_ragged"] = _sequence("C2", 5, 5, VS_RANSAC_DEVICE, fb=1)
    res["seq_lanes"] = _sequence("C2", 6, 4, VS_RANSAC_OFF, lanes=3)
    res["seq_odd"] = _sequence("C1", 2, 4, VS_RANSAC_OFF, sizes=(333, 251))
    res["seq_small"] = _sequence("C1", 2, 3, VS_RANSAC_OFF, sizes=(97, 96))
    res["stages_c1"] = _stages("C1")
    res["stages_c3"] = _stages("C3")
    sys.stdout.buffer.write(pickle.dumps(res))


def _run(guard):
    env = dict(os.environ)
    env.pop("VS_DEBUG_GUARD", None)
    if guard:
        env["VS_DEBUG_GUARD"] = "1"
    code = f"import sys; sys.path.insert(0, {ROOT!r}); sys.path.insert(0, {os.path.join(ROOT, 'tests')!r}); " \
           f"import test_gpu_guard as T; T._child()"
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, cwd=ROOT)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    import pickle
    return pickle.loads(r.stdout)


def _same(a, b):
    if isinstance(a, (list, tuple)):
        return len(a) == len(b) and all(_same(x, y) for x, y in zip(a, b))
    return np.array_equal(np.asarray(a), np.asarray(b), equal_nan=True)


def test_no_store_leaves_its_buffer(built_lib):
    guarded, plain = _run(True), _run(False)
    for key, (out, g) in guarded.items():
        assert g == 0, f"{key}: {g} guard bytes overwritten"
        assert plain[key][1] == -1, "an unguarded handle must say so"
        assert _same(out, plain[key][0]), key      # the device RANSAC seeds itself from the stream's points: deterministic
