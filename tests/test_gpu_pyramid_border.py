"""The LK kernel reads its 48-byte tile rows straight out of the stored levels, border included: every stored level must
carry cv::buildOpticalFlowPyramid's REFLECT_101 border (48 columns, 32 rows here) around the exact level — checked byte by
byte on the frame path's own pyramid (vs_tracker_debug_read_pyramid), for shapes that take the in-producer border path and
shapes that fall back to the border kernels (odd widths, narrow levels), with and without CLAHE."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from mobile_slam_b200 import BatchTracker, VS_RANSAC_OFF, config_from_synth  # noqa: E402
from mobile_slam_b200.synth import CONFIGS  # noqa: E402

pytestmark = pytest.mark.gpu

BASE = dict(max_cnt=60, min_dist=12, max_level=3, fisheye=0, model="PINHOLE", dist=(0.0, 0.0, 0.0, 0.0), focal_length=460.0)
SHAPES = [(752, 480), (512, 512), (640, 480), (333, 251), (336, 250), (200, 131), (97, 96), (1241, 376), (520, 300),
          (128, 66), (136, 70), (72, 72)]


@pytest.mark.parametrize("w,h", SHAPES)
@pytest.mark.parametrize("equalize", [1, 0])
def test_stored_levels_carry_reflect101_border(built_lib, w, h, equalize):
    cfg = dict(BASE, width=w, height=h, intr=(0.8 * w, 0.8 * w, w / 2, h / 2))
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF, equalize=equalize)
    n_streams = 3
    bt = BatchTracker(tc, cam, n_streams)
    rng = np.random.default_rng(w * 1000 + h)
    for k in range(2):                                  # both pyramid buffers
        frames = [rng.integers(0, 256, (h, w), dtype=np.uint8) for _ in range(n_streams)]
        bt.track(frames, [0.05 * k] * n_streams)
        for s in (0, n_streams - 1):
            l0 = cv2.createCLAHE(3.0, (8, 8)).apply(frames[s]) if equalize else frames[s]
            _, pyr = cv2.buildOpticalFlowPyramid(l0, (21, 21), 3, withDerivatives=False)
            got = bt.debug_read_pyramid(s)
            assert len(got) == len(pyr)
            for l, ((blk, px, py), lev) in enumerate(zip(got, pyr)):
                # REFLECT_101 needs pad < size; narrower levels are only required to hold the border a reflection defines
                ex, ey = min(px, lev.shape[1] - 1), min(py, lev.shape[0] - 1)
                want = np.pad(lev, ((ey, ey), (ex, ex)), mode="reflect")
                have = blk[py - ey:py + lev.shape[0] + ey, px - ex:px + lev.shape[1] + ex]
                bad = np.argwhere(have != want)
                assert bad.size == 0, f"{w}x{h} eq={equalize} frame {k} stream {s} level {l}: {len(bad)} bytes differ, first at (row, col) " \
                                      f"{tuple(bad[0] - (ey, ex))}"
    bt.close()
