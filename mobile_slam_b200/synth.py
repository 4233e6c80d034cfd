"""Seeded synthetic textured camera streams (SURVEY.md §8(d) recipe).

Real TUM-VI / EuRoC / KITTI-360 frames are absent (git-ignored upstream, no network), so every test
and benchmark runs on sequences of the same *shape*: a band-limited random texture on a 2x canvas,
viewed through a smooth similarity warp (rotation <= 2 deg/frame, translation <= 4 px/frame, scale
1 +- 0.002), plus sigma=1 pixel noise and a slow gain drift.  Seeds are ``stream_id*1000 + frame``.

Pure numpy/scipy so the same frames can be produced here, on the GPU box and inside the oracle's
fixtures without OpenCV.
"""
from __future__ import annotations

import numpy as np
from scipy import ndimage

__all__ = ["SynthStream", "make_frames", "CONFIGS"]

# BASELINE.json configs -> (width, height, max_cnt, min_dist, max_level, camera)
CONFIGS = {
    # TUM-VI room1 (config/tum_vi_room1.yaml:14-52): equidistant, fisheye mask on
    "C1": dict(width=512, height=512, max_cnt=150, min_dist=20, max_level=3, fisheye=1,
               model="KANNALA_BRANDT",
               intr=(190.9784771512872, 190.9733070521226, 254.9317060593548, 256.8974428996504),
               dist=(0.0034823894022493434, 0.0007150348452162257, -0.0020532361418706202,
                     0.00020293673591811182), focal_length=460.0),
    # EuRoC MH_01 (config/euroc.yaml:14-50): pinhole radtan
    "C2": dict(width=752, height=480, max_cnt=150, min_dist=30, max_level=3, fisheye=0,
               model="PINHOLE", intr=(461.6, 460.3, 363.0, 248.1),
               dist=(-0.2917, 0.08228, 5.333e-05, -1.578e-04), focal_length=460.0),
    # KITTI-360 (config/kitti360.yaml:11-48): pinhole, no distortion; BASELINE says 400 features
    "C3": dict(width=1408, height=376, max_cnt=400, min_dist=30, max_level=4, fisheye=0,
               model="PINHOLE", intr=(552.554261, 552.554261, 682.049453, 238.769549),
               dist=(0.0, 0.0, 0.0, 0.0), focal_length=460.0),
    # synthetic feature-heavy stress
    "C4": dict(width=1280, height=720, max_cnt=2000, min_dist=10, max_level=4, fisheye=0,
               model="PINHOLE", intr=(700.0, 700.0, 640.0, 360.0),
               dist=(-0.05, 0.01, 1e-4, -1e-4), focal_length=460.0),
}


class SynthStream:
    """One independent synthetic camera stream."""

    def __init__(self, stream_id: int, width: int, height: int, *, motion: float = 1.0,
                 noise_sigma: float = 1.0, gain_drift: float = 0.04):
        self.stream_id = int(stream_id)
        self.w, self.h = int(width), int(height)
        self.noise_sigma = float(noise_sigma)
        self.gain_drift = float(gain_drift)
        rng = np.random.default_rng(self.stream_id * 1000 + 999)
        ch, cw = 2 * self.h, 2 * self.w
        a = ndimage.gaussian_filter(rng.random((ch, cw), dtype=np.float32), 1.5, mode="wrap")
        b = ndimage.gaussian_filter(rng.random((ch, cw), dtype=np.float32), 6.0, mode="wrap")
        a = (a - a.mean()) / (a.std() + 1e-9)
        b = (b - b.mean()) / (b.std() + 1e-9)
        t = 0.6 * a + 0.4 * b
        lo, hi = np.percentile(t, 0.5), np.percentile(t, 99.5)
        self.canvas = np.clip((t - lo) / (hi - lo), 0.0, 1.0).astype(np.float32) * 255.0
        # smooth motion parameters (per-frame rates bounded as in the recipe)
        m = float(motion)
        self.rot_amp = np.deg2rad(rng.uniform(4.0, 10.0))           # total swing
        self.rot_period = rng.uniform(60.0, 140.0) / max(m, 1e-6)   # => <= ~1 deg/frame * m
        self.tx_amp = rng.uniform(40.0, 0.20 * self.w)
        self.ty_amp = rng.uniform(40.0, 0.20 * self.h)
        self.tx_period = max(2 * np.pi * self.tx_amp / (3.5 * m), 40.0)
        self.ty_period = max(2 * np.pi * self.ty_amp / (3.5 * m), 40.0)
        self.ph = rng.uniform(0, 2 * np.pi, size=4)
        self.scale_amp = 0.03
        self.scale_period = rng.uniform(150.0, 300.0) / max(m, 1e-6)
        yy, xx = np.mgrid[0:self.h, 0:self.w].astype(np.float32)
        self._xx = xx - (self.w - 1) * 0.5
        self._yy = yy - (self.h - 1) * 0.5

    def pose(self, k: int):
        """(theta, scale, tx, ty) of frame k."""
        th = self.rot_amp * np.sin(2 * np.pi * k / self.rot_period + self.ph[0])
        s = 1.0 + self.scale_amp * np.sin(2 * np.pi * k / self.scale_period + self.ph[1])
        tx = self.tx_amp * np.sin(2 * np.pi * k / self.tx_period + self.ph[2])
        ty = self.ty_amp * np.sin(2 * np.pi * k / self.ty_period + self.ph[3])
        return float(th), float(s), float(tx), float(ty)

    def warp_point(self, k: int, x, y):
        """Canvas coordinate seen by frame-k pixel (x, y) (for ground-truth flow in tests)."""
        th, s, tx, ty = self.pose(k)
        c, sn = np.cos(th) * s, np.sin(th) * s
        xc = np.asarray(x, dtype=np.float64) - (self.w - 1) * 0.5
        yc = np.asarray(y, dtype=np.float64) - (self.h - 1) * 0.5
        return (c * xc - sn * yc + tx + (2 * self.w - 1) * 0.5,
                sn * xc + c * yc + ty + (2 * self.h - 1) * 0.5)

    def frame(self, k: int) -> np.ndarray:
        th, s, tx, ty = self.pose(k)
        c, sn = np.float32(np.cos(th) * s), np.float32(np.sin(th) * s)
        cx = c * self._xx - sn * self._yy + np.float32(tx + (2 * self.w - 1) * 0.5)
        cy = sn * self._xx + c * self._yy + np.float32(ty + (2 * self.h - 1) * 0.5)
        img = ndimage.map_coordinates(self.canvas, [cy, cx], order=1, mode="mirror")
        gain = 1.0 + self.gain_drift * np.sin(k / 37.0 + self.ph[0])
        img = img * np.float32(gain)
        if self.noise_sigma > 0:
            rng = np.random.default_rng(self.stream_id * 1000 + k)
            img = img + rng.normal(0.0, self.noise_sigma, img.shape).astype(np.float32)
        return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def make_frames(stream_id: int, width: int, height: int, n_frames: int, **kw) -> np.ndarray:
    """(n_frames, H, W) uint8 frames of one stream."""
    s = SynthStream(stream_id, width, height, **kw)
    return np.stack([s.frame(k) for k in range(n_frames)])
