"""Parity sweeps (scripts/shape_sweep.py, scripts/param_sweep.py) as tests: 33 image shapes (odd widths / heights, very
wide, very narrow, tile-grid aligned on one side only) through every stage and a short sequence, and 45 parameter
combinations (LK windows 5..31, 0..5 pyramid levels, min_dist 1..200, max_cnt 1..1000, iteration / eps criteria,
equalize off, fisheye on/off) through 6-frame sequences, all against OpenCV / the reference restatement."""
import os
import sys

import pytest

pytest.importorskip("cv2")
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))

pytestmark = pytest.mark.gpu


def test_shape_sweep(built_lib):
    from shape_sweep import shape_sweep
    bad = shape_sweep(n_random=12, verbose=False)
    assert not bad, bad


def test_param_sweep(built_lib):
    from param_sweep import param_sweep
    bad = param_sweep(verbose=False)
    assert not bad, bad


@pytest.mark.parametrize("name,motion,frames", [("C1", 1.0, 300), ("C2", 3.0, 300), ("C3", 2.0, 120)])
def test_long_free_running_sequences(built_lib, name, motion, frames):
    # 300 frames = the length of the reference's own trajectory parity test (tests/test_vio_engine_parity.cpp:305); C2 runs
    # with 3x motion (EuRoC's frame_skip 2).  A free-running pair may renumber features after one
    # borderline cvRound / status flip (1e-4 px LK differences), so ids are compared up to the first divergence and
    # image positions over the whole run.
    from long_sequence import long_sequence
    r = long_sequence(name, motion, frames, verbose=False)
    assert r["first_divergence"] is None or r["first_divergence"] >= 60, r
    assert r["max_dpos_before_divergence"] <= 0.011, r
    assert r["pos_match_frac"] >= 0.97, r


def test_camera_calibration_sweep(built_lib):
    # 42 calibrations over the four camera models (pinhole incl. zero / one-sided distortion, Kannala-Brandt with every
    # trailing-zero pattern, Mei below / at / above xi = 1, Scaramuzza), pixels inside and 40 px outside the image:
    # float32-identical to the fp64 restatement (<= 1 ulp for Kannala-Brandt), NaN where the reference gives NaN
    from camera_sweep import camera_sweep
    bad = camera_sweep(verbose=False)
    assert not bad, bad


def test_soak_submit_wait_and_handle_lifecycle(built_lib):
    # 600 batches through submit/wait on a 2-lane handle with the device RANSAC, then 60 create/track/destroy cycles:
    # no error, no collapse of the published feature count, no device-memory growth (scripts/soak.py asserts them)
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "scripts", "soak.py"), "600", "8"], capture_output=True, text=True,
                       timeout=600, cwd=root)
    assert r.returncode == 0 and "soak ok" in r.stdout, r.stdout[-800:] + r.stderr[-800:]
