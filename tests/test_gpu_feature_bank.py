"""The device feature bank (vs_bank_*, csrc/feature_bank.cu) against the FeatureManager oracle (oracle/feature_manager.py,
itself bit-equal to the reference's compiled feature_manager.cpp — tests/test_oracle_feature_manager.py), through the
C ABI: SURVEY.md §8 f4, /root/reference/src/frontend/feature_manager.cpp:27-73,142-294.

Bars: key-frame decisions, last_track_num, list order, ids, start frames, observation counts, solve flags and every stored
observation identical; parallax sums bit-equal; depths within 1e-9 relative (the triangulation's SVD is a Jacobi iteration
on the device and LAPACK behind the oracle; Eigen's own JacobiSVD is a third)."""
import numpy as np
import pytest

from mobile_slam_b200 import BatchTracker, VsError, VS_RANSAC_OFF, config_from_synth
from mobile_slam_b200.feature_bank import (OP_NONE, OP_REMOVE_BACK, OP_REMOVE_BACK_SHIFT_DEPTH, OP_REMOVE_FRONT, WINDOW_SIZE,
                                           FeatureBank, StreamFeatureManager)
from mobile_slam_b200.synth import CONFIGS, SynthStream
from oracle import fm_scenario
from oracle.feature_manager import FeatureManager

pytestmark = pytest.mark.gpu
DEPTH_RTOL = 1e-9


def _same_state(dev, ora, where, depth_rtol=DEPTH_RTOL):
    ids, start, n_obs, depth, flag, obs = dev.export()
    oi, os_, on, od, of, oo = ora.export()
    assert len(ids) == len(oi), f"{where}: bank size {len(ids)} vs {len(oi)}"
    assert np.array_equal(ids, oi), f"{where}: ids / list order"
    assert np.array_equal(start, os_), f"{where}: start_frame"
    assert np.array_equal(n_obs, on), f"{where}: observation counts"
    assert np.array_equal(flag, of), f"{where}: solve_flag"
    assert np.array_equal(obs.view(np.uint8), oo.view(np.uint8)), f"{where}: observations"
    assert np.allclose(depth, od, rtol=depth_rtol, atol=0), f"{where}: depth, max rel {np.max(np.abs(depth - od) / np.abs(od))}"
    return float(np.max(np.abs(depth - od) / np.abs(od))) if len(od) else 0.0


def _lockstep(dev, ora, bank, stream, seed, n_frames, **kw):
    gen = fm_scenario.steps(seed, n_frames, **kw)
    op = next(gen, None)
    n_ops, keys, worst = {}, [], 0.0
    while op is not None:
        n_ops[op[0]] = n_ops.get(op[0], 0) + 1
        send = None
        if op[0] == "add":
            k1 = ora.add_feature_and_check_parallax(op[1], op[2])
            k2 = dev.add_feature_and_check_parallax(op[1], op[2])
            assert k1 == k2, f"key-frame decision at add #{n_ops['add']}"
            assert ora.last_track_num == dev.last_track_num
            ps, pn = ora.last_parallax
            assert pn == bank.parallax_num[stream]
            assert np.float64(ps).tobytes() == np.float64(bank.parallax_sum[stream]).tobytes(), "parallax sum not bit-equal"
            keys.append(k1)
            send = k1
        elif op[0] == "triangulate":
            ora.triangulate_across_all_views(op[1], op[2], op[3])
            dev.triangulate_across_all_views(op[1], op[2], op[3])
        elif op[0] == "depth_vector":
            d1, d2 = ora.get_depth_vector(), dev.get_depth_vector()
            assert len(d1) == len(d2) == dev.get_feature_count() == ora.get_feature_count()
            assert np.allclose(d1, d2, rtol=DEPTH_RTOL, atol=0)
            send = d1
        elif op[0] == "set_depth":
            ora.set_depth(op[1]); dev.set_depth(op[1])
        elif op[0] == "remove_failures":
            ora.remove_failures(); dev.remove_failures()
        elif op[0] == "remove_back":
            ora.remove_back(); dev.remove_back()
        elif op[0] == "remove_back_shift_depth":
            ora.remove_back_shift_depth(*op[1:]); dev.remove_back_shift_depth(*op[1:])
        elif op[0] == "remove_front":
            ora.remove_front(op[1]); dev.remove_front(op[1])
        worst = max(worst, _same_state(dev, ora, f"seed {seed} after {op[0]} #{n_ops[op[0]]}"))
        try:
            op = gen.send(send)
        except StopIteration:
            op = None
    c1, c2 = ora.get_corresponding(WINDOW_SIZE - 3, WINDOW_SIZE), dev.get_corresponding(WINDOW_SIZE - 3, WINDOW_SIZE)
    assert c1 == c2
    return keys, n_ops, worst


@pytest.mark.parametrize("seed,speed", [(1, 0.06), (2, 0.004), (3, 0.02), (4, 0.15)])
def test_bank_equals_oracle_over_estimator_loop(built_lib, seed, speed):
    bank = FeatureBank(3, capacity=1024)            # the scenario runs on stream 1; streams 0 and 2 must stay empty
    dev, ora = StreamFeatureManager(bank, 1), FeatureManager(10.0, 460.0, 5.0)
    keys, n_ops, worst = _lockstep(dev, ora, bank, 1, seed, 60, speed=speed)
    assert n_ops["add"] == 60 and n_ops.get("triangulate", 0) > 30
    assert len(bank.export(0)[0]) == 0 and len(bank.export(2)[0]) == 0
    print(f"seed {seed}: {sum(keys)} key frames of 60, ops {n_ops}, worst relative depth difference {worst:.2e}")
    bank.close()


def test_bank_batch_of_streams_equals_single_streams(built_lib):
    """8 streams with different scenarios advance in ONE launch per operation; every stream equals its own oracle."""
    B = 8
    bank = FeatureBank(B, capacity=1024)
    oras = [FeatureManager(10.0, 460.0, 5.0) for _ in range(B)]
    scs = [fm_scenario.Scenario(100 + s, speed=[0.004, 0.02, 0.06, 0.15][s % 4]) for s in range(B)]
    fcs = np.zeros(B, np.int32)
    windows = [[None] * (WINDOW_SIZE + 1) for _ in range(B)]
    nonlinear = [False] * B
    for k in range(40):
        images, poses = [], []
        for s in range(B):
            im, pose = scs[s].frame()
            images.append(im)
            windows[s][fcs[s]] = pose
        keys = bank.add_frames(fcs, images)
        for s in range(B):
            assert bool(keys[s]) == oras[s].add_feature_and_check_parallax(int(fcs[s]), images[s]), f"stream {s} frame {k}"
            assert bank.last_track_num[s] == oras[s].last_track_num
        full = fcs == WINDOW_SIZE
        tri = []
        for s in range(B):
            if full[s] and nonlinear[s]:
                w = [p if p is not None else windows[s][fcs[s]] for p in windows[s]]
                tri.append(w)
                oras[s].triangulate_across_all_views(w, scs[s].t_ic, scs[s].r_ic)
            else:
                tri.append(None)
        if any(t is not None for t in tri):
            assert all(np.array_equal(scs[0].t_ic, sc.t_ic) for sc in scs)
            bank.triangulate(tri, scs[0].t_ic, scs[0].r_ic)
        ops, rp = [], []
        for s in range(B):
            if not full[s]:
                ops.append(OP_NONE); rp.append(None)
                fcs[s] += 1
                continue
            sc, win = scs[s], windows[s]
            if keys[s]:
                back_R0, back_P0 = win[0]
                win[:] = win[1:] + [win[WINDOW_SIZE]]
                if nonlinear[s]:
                    q = (back_R0 @ sc.r_ic, back_P0 + back_R0 @ sc.t_ic, win[0][0] @ sc.r_ic, win[0][1] + win[0][0] @ sc.t_ic)
                    oras[s].remove_back_shift_depth(*q)
                    ops.append(OP_REMOVE_BACK_SHIFT_DEPTH); rp.append(q)
                else:
                    oras[s].remove_back()
                    ops.append(OP_REMOVE_BACK); rp.append(None)
            else:
                win[WINDOW_SIZE - 1] = win[WINDOW_SIZE]
                oras[s].remove_front(int(fcs[s]))
                ops.append(OP_REMOVE_FRONT); rp.append(None)
            nonlinear[s] = True
        bank.slide(ops, fcs, rp)
        for s in range(B):
            _same_state(StreamFeatureManager(bank, s), oras[s], f"stream {s} frame {k}")
    assert len({len(o.bank) for o in oras}) > 1          # the streams really went different ways
    bank.close()


@pytest.mark.parametrize("n_streams", [2, 20])
def test_bank_fed_from_the_tracker_on_the_device(built_lib, n_streams):
    """Tracker results never leave the GPU on their way into the bank: decisions and bank content equal the oracle fed with
    the records the reference's publish loop builds from the same results (2 streams: pinned result slots; 20: device block)."""
    cfg = CONFIGS["C2"]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    bt = BatchTracker(tc, cam, n_streams)
    bank = FeatureBank(n_streams, capacity=2048)
    oras = [FeatureManager(10.0, 460.0, 5.0) for _ in range(n_streams)]
    streams = [SynthStream(300 + s, cfg["width"], cfg["height"]) for s in range(n_streams)]
    fcs = np.zeros(n_streams, np.int32)
    n_key = 0
    for k in range(16):
        res = bt.track([st.frame(k) for st in streams], [0.05 * k] * n_streams)
        keys = bank.add_frames_from_tracker(bt, fcs)
        for s, r in enumerate(res):
            image = {}
            for j in range(len(r.ids)):
                if r.track_cnt[j] > 1:      # measurement_processor.cpp:210-231
                    image[int(r.ids[j])] = [float(r.cur_undistorted_pts[j, 0]), float(r.cur_undistorted_pts[j, 1]), 1.0,
                                            float(r.cur_pts[j, 0]), float(r.cur_pts[j, 1]),
                                            float(r.pts_velocity[j, 0]), float(r.pts_velocity[j, 1])]
            assert bool(keys[s]) == oras[s].add_feature_and_check_parallax(int(fcs[s]), image), f"stream {s} frame {k}"
            assert bank.last_track_num[s] == oras[s].last_track_num
            n_key += int(keys[s])
        ops = []
        for s in range(n_streams):
            if fcs[s] < WINDOW_SIZE:
                fcs[s] += 1
                ops.append(OP_NONE)
            elif keys[s]:
                oras[s].remove_back(); ops.append(OP_REMOVE_BACK)
            else:
                oras[s].remove_front(int(fcs[s])); ops.append(OP_REMOVE_FRONT)
        bank.slide(ops, fcs)
        for s in (0, n_streams - 1):
            _same_state(StreamFeatureManager(bank, s), oras[s], f"stream {s} frame {k}")
    assert len(oras[0].bank) > 100 and n_key > 0
    bank.close(); bt.close()


def test_bank_errors(built_lib):
    bank = FeatureBank(1, capacity=8)
    img = {i: [0.01 * i, 0.0, 1.0, 0, 0, 0, 0] for i in range(9)}
    with pytest.raises(VsError, match="full"):
        bank.add_frames(0, [img])
    bank.close()
    with pytest.raises(VsError):
        FeatureBank(0)
    bank = FeatureBank(2, capacity=16)
    with pytest.raises(VsError, match="removeFront needs frame_count"):
        bank.slide([OP_REMOVE_FRONT, OP_NONE])
    with pytest.raises(VsError, match="out of range"):
        bank.export(2)
    bank.close()


def _guard_child():
    """Runs in a child interpreter with VS_DEBUG_GUARD=1: every bank kernel on small and on overflowing inputs, then the
    number of guard bytes that changed around the bank's device blocks."""
    import sys
    worst = 0
    bank = FeatureBank(3, capacity=300)
    for s, (seed, speed) in enumerate([(1, 0.06), (2, 0.004), (4, 0.15)]):
        fm_scenario.drive(StreamFeatureManager(bank, s), seed, 40, speed=speed)
    worst = max(worst, bank.guard_check())
    bank.close()
    bank = FeatureBank(2, capacity=8)                      # more features than slots: refused, nothing written past the end
    img = {i: [0.01 * i, 0.0, 1.0, 0, 0, 0, 0] for i in range(40)}
    for k in range(3):
        try:
            bank.add_frames(k, [img, img])
        except VsError:
            pass
    worst = max(worst, bank.guard_check())
    bank.close()
    cfg = CONFIGS["C1"]
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    bt = BatchTracker(tc, cam, 2)
    bank = FeatureBank(2, capacity=160)                    # fills up within the window: the overflow path with tracker input
    st = [SynthStream(5 + s, cfg["width"], cfg["height"], motion=3.0) for s in range(2)]
    for k in range(14):
        bt.track([q.frame(k) for q in st], [0.05 * k] * 2)
        try:
            bank.add_frames_from_tracker(bt, min(k, WINDOW_SIZE))
        except VsError:
            pass
    worst = max(worst, bank.guard_check())
    bank.close(); bt.close()
    sys.stdout.write(f"GUARD {worst}\n")


def test_bank_kernels_store_inside_their_buffers(built_lib):
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (f"import sys; sys.path.insert(0, {root!r}); sys.path.insert(0, {os.path.join(root, 'tests')!r}); "
            "import test_gpu_feature_bank as T; T._guard_child()")
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, VS_DEBUG_GUARD="1"), capture_output=True, text=True,
                       cwd=root, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("GUARD ")][-1]
    assert int(line.split()[1]) == 0, f"{line}: guard bytes around the bank's device blocks were overwritten"
    bank = FeatureBank(1, capacity=8)
    assert bank.guard_check() == -1                        # created without guards: says so
    bank.close()


def test_erased_feature_returns_at_the_back_of_the_list(built_lib):
    """removeFailures erases a feature the tracker still publishes: its id comes back as a NEW entry at the back, so the list
    is no longer ordered by id — the device bank must keep the reference's list order (pinned against the compiled reference
    in tests/test_oracle_feature_manager.py)."""
    from test_oracle_feature_manager import _apply_reuse, _reuse_ops
    bank = FeatureBank(2, capacity=64)
    dev, ora = StreamFeatureManager(bank, 1), FeatureManager()
    for k, op in enumerate(_reuse_ops()):
        r1, r2 = _apply_reuse(ora, op), _apply_reuse(dev, op)
        assert r1 == r2, (k, op[0])
        _same_state(dev, ora, f"reuse op {k} {op[0]}")
    ids = list(dev.export()[0])
    assert ids != sorted(ids) and ids.index(5) > ids.index(30)
    assert len(bank.export(0)[0]) == 0
    bank.close()
