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
