"""TEST INFRASTRUCTURE ONLY — the reference's OpenCV CPU tracker, restated line by line.

``RefFeatureTracker`` follows ``/root/reference/src/frontend/feature_tracker.cpp`` statement by
statement and makes the same five OpenCV calls through cv2 (the reference cannot be compiled here:
it needs OpenCV C++/Eigen/Ceres, none installed — DESIGN.md "Oracle").  Used (a) as the parity
checker in tests/, (b) as the timed CPU baseline in bench.py.  Never imported by the product.

Reference lines are quoted as ``ft:NNN``.
"""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass, field

import cv2
import numpy as np

from . import camera as cam

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def _stdsort_lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "_build", "libstdsort.so")
        if not os.path.exists(path):
            import subprocess
            subprocess.check_call(["make", "-C", _HERE, "-s"])
        _LIB = ctypes.CDLL(path)
        _LIB.oracle_setmask_order.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    return _LIB


def setmask_order(track_cnt) -> np.ndarray:
    """Permutation produced by the reference's unstable ``std::sort`` (ft:70-73), real libstdc++."""
    tc = np.ascontiguousarray(track_cnt, np.int32)
    perm = np.zeros(len(tc), np.int32)
    if len(tc):
        _stdsort_lib().oracle_setmask_order(tc.ctypes.data, len(tc), perm.ctypes.data)
    return perm


@dataclass
class RefConfig:
    """utility::g_config fields the tracker reads (include/utility/config.h:18-56)."""
    row: int = 480
    col: int = 752
    focal_length: float = 460.0
    max_cnt: int = 150
    min_dist: int = 30
    f_threshold: float = 1.0
    equalize: int = 1
    fisheye: int = 0
    lk_window_size: int = 21
    lk_pyramid_levels: int = 3
    lk_iterations: int = 30
    lk_eps: float = 0.01
    f_threshold_edge_factor: float = 0.0
    use_ransac: bool = True          # False = skip rejectWithFundamentalMatrix (stage isolation)


def _cv_round(v) -> int:
    return int(np.rint(v))


class RefFeatureTracker:
    n_id = 0  # static int FeatureTracker::n_id (ft:10) — reset per tracker via reset_ids()

    def __init__(self, cfg: RefConfig, camera: cam.Camera, per_instance_ids: bool = True):
        self.cfg = cfg
        self.m_camera = camera
        self.per_instance_ids = per_instance_ids
        self._n_id = 0
        self.mask = None
        self.fisheye_mask = None
        self.cur_img = None
        self.next_img = None
        self.n_pts = np.zeros((0, 2), np.float32)
        self.prev_pts = np.zeros((0, 2), np.float32)
        self.cur_pts = np.zeros((0, 2), np.float32)
        self.next_pts = np.zeros((0, 2), np.float32)
        self.cur_undistorted_pts = np.zeros((0, 2), np.float32)
        self.pts_velocity = np.zeros((0, 2), np.float32)
        self.ids: list[int] = []
        self.track_cnt: list[int] = []
        self.cur_undistorted_pts_map: dict[int, tuple] = {}
        self.prev_undistorted_pts_map: dict[int, tuple] = {}
        self.cur_time = 0.0
        self.prev_time = 0.0
        self.clahe_ = None
        self.stage = {}          # per-stage dumps of the last call (for teacher-forced parity tests)
        self.timing = {}

    # ft:12-18
    def in_border(self, pt) -> bool:
        x, y = _cv_round(pt[0]), _cv_round(pt[1])
        return 1 <= x < self.cfg.col - 1 and 1 <= y < self.cfg.row - 1

    # ft:38-87
    def set_mask(self):
        c = self.cfg
        if c.fisheye:
            if self.fisheye_mask is None:
                fm = np.zeros((c.row, c.col), np.uint8)
                cx, cy = c.col / 2.0, c.row / 2.0
                radius = min(cx, cy, c.col - cx, c.row - cy) * 0.95
                cv2.circle(fm, (int(cx), int(cy)), int(radius), 255, -1)
                self.fisheye_mask = fm
            self.mask = self.fisheye_mask.copy()
        else:
            self.mask = np.full((c.row, c.col), 255, np.uint8)
        perm = setmask_order(self.track_cnt)
        pts, ids, cnt = self.next_pts, self.ids, self.track_cnt
        kp, ki, kc = [], [], []
        for j in perm:
            x, y = _cv_round(pts[j][0]), _cv_round(pts[j][1])
            if self.mask[y, x] == 255:
                kp.append(pts[j]); ki.append(ids[j]); kc.append(cnt[j])
                cv2.circle(self.mask, (x, y), c.min_dist, 0, -1)
        self.next_pts = np.array(kp, np.float32).reshape(-1, 2)
        self.ids, self.track_cnt = ki, kc

    def _filter(self, status):
        keep = np.asarray(status, bool)
        self.prev_pts = self.prev_pts[keep] if len(self.prev_pts) == len(keep) else self.prev_pts
        self.cur_pts = self.cur_pts[keep]
        self.next_pts = self.next_pts[keep]
        self.cur_undistorted_pts = self.cur_undistorted_pts[keep]
        self.ids = [v for v, k in zip(self.ids, keep) if k]
        self.track_cnt = [v for v, k in zip(self.track_cnt, keep) if k]

    # ft:97-201
    def detect_and_track(self, img: np.ndarray, t: float):
        c = self.cfg
        self.cur_time = t
        st = self.stage = {}
        if c.equalize:
            if self.clahe_ is None:
                self.clahe_ = cv2.createCLAHE(3.0, (8, 8))
            img = self.clahe_.apply(img)
        st["clahe"] = img
        win = (c.lk_window_size, c.lk_window_size)
        first = self.next_img is None
        self.next_img = img
        if first:
            self.cur_img = img
        # ft:123 — the pyramid is built inside calcOpticalFlowPyrLK from the images (cv2's Python
        # binding rejects pyramid lists); identical arithmetic.
        self.next_pts = np.zeros((0, 2), np.float32)
        st["lk_in"] = self.cur_pts.copy()
        if len(self.cur_pts) > 0:
            crit = (cv2.TERM_CRITERIA_COUNT | cv2.TERM_CRITERIA_EPS, c.lk_iterations, c.lk_eps)
            nxt, status, _err = cv2.calcOpticalFlowPyrLK(
                self.cur_img, self.next_img, self.cur_pts.reshape(-1, 1, 2), None,
                winSize=win, maxLevel=c.lk_pyramid_levels, criteria=crit)
            nxt = nxt.reshape(-1, 2)
            status = status.reshape(-1).astype(np.uint8)
            st["lk_out"], st["lk_status_raw"] = nxt.copy(), status.copy()
            for i in range(len(nxt)):
                if status[i] and not self.in_border(nxt[i]):
                    status[i] = 0
            st["lk_status"] = status.copy()
            self.next_pts = nxt
            self._filter(status)
        self.track_cnt = [n + 1 for n in self.track_cnt]
        # ft:156-173
        should_reject = True
        if len(self.next_pts) > 0 and len(self.cur_pts) == len(self.next_pts):
            # ft:161-162: `double dx = next_pts[i].x - cur_pts[i].x;` — a float32 subtraction, widened afterwards
            d = (self.next_pts.astype(np.float32) - self.cur_pts.astype(np.float32)).astype(np.float64)
            rms = np.sqrt((d * d).sum() / len(self.next_pts))
            if rms < 2.0:
                should_reject = False
        st["ransac_in"] = (self.cur_pts.copy(), self.next_pts.copy())
        st["ransac_ran"] = False
        if should_reject and c.use_ransac:
            self.reject_with_f()
        st["pre_mask"] = (self.next_pts.copy(), list(self.ids), list(self.track_cnt))
        self.set_mask()
        st["post_mask"] = (self.next_pts.copy(), list(self.ids), list(self.track_cnt))
        st["mask"] = self.mask
        n_new = c.max_cnt - len(self.next_pts)
        if n_new > 0:
            r = cv2.goodFeaturesToTrack(self.next_img, n_new, 0.01, c.min_dist, mask=self.mask)
            self.n_pts = np.zeros((0, 2), np.float32) if r is None else r.reshape(-1, 2)
        else:
            self.n_pts = np.zeros((0, 2), np.float32)
        st["n_pts"] = self.n_pts.copy()
        # addPoints ft:89-95
        self.next_pts = np.concatenate([self.next_pts.reshape(-1, 2), self.n_pts]).astype(np.float32)
        self.ids += [-1] * len(self.n_pts)
        self.track_cnt += [1] * len(self.n_pts)
        self.prev_pts = self.cur_pts
        self.cur_img = self.next_img
        self.cur_pts = self.next_pts
        self.undistorted_points()
        self.prev_time = self.cur_time

    # ft:203-307
    def reject_with_f(self):
        c = self.cfg
        if len(self.next_pts) < 30:
            return
        cx, cy = c.col / 2.0, c.row / 2.0

        def lift(pts):
            P = self.m_camera.lift_batch(pts)
            return np.stack([c.focal_length * P[:, 0] / P[:, 2] + cx,
                             c.focal_length * P[:, 1] / P[:, 2] + cy], 1).astype(np.float32)

        un_cur, un_next = lift(self.cur_pts), lift(self.next_pts)
        F, status = cv2.findFundamentalMat(un_cur, un_next, cv2.FM_RANSAC, c.f_threshold, 0.99)
        self.stage["ransac_ran"] = True
        self.stage["ransac_pts"] = (un_cur, un_next)
        if status is None:
            status = np.zeros(len(un_cur), np.uint8)
        status = status.reshape(-1).astype(np.uint8)
        status = edge_recovery(status, F, un_cur, un_next, self.next_pts, c)
        self.stage["ransac_status"] = status.copy()
        self._filter(status)

    # ft:309-316
    def update_id(self, i: int) -> bool:
        if i < len(self.ids):
            if self.ids[i] == -1:
                if self.per_instance_ids:
                    self.ids[i] = self._n_id
                    self._n_id += 1
                else:
                    self.ids[i] = RefFeatureTracker.n_id
                    RefFeatureTracker.n_id += 1
            return True
        return False

    def update_all_ids(self):
        i = 0
        while self.update_id(i):
            i += 1

    # ft:338-380
    def undistorted_points(self):
        self.cur_undistorted_pts = self.m_camera.undistort_f32(self.cur_pts)
        self.cur_undistorted_pts_map = {}
        for i in range(len(self.cur_pts)):
            self.cur_undistorted_pts_map.setdefault(self.ids[i], tuple(self.cur_undistorted_pts[i]))
        n = len(self.cur_pts)
        vel = np.zeros((n, 2), np.float32)
        if self.prev_undistorted_pts_map:
            dt = self.cur_time - self.prev_time
            for i in range(n):
                if self.ids[i] != -1 and self.ids[i] in self.prev_undistorted_pts_map and dt > 1e-4:
                    px, py = self.prev_undistorted_pts_map[self.ids[i]]
                    # float - float is a float32 subtraction in the reference, then / double dt
                    vel[i, 0] = np.float32(float(np.float32(self.cur_undistorted_pts[i][0]) - np.float32(px)) / dt)
                    vel[i, 1] = np.float32(float(np.float32(self.cur_undistorted_pts[i][1]) - np.float32(py)) / dt)
        self.pts_velocity = vel
        self.prev_undistorted_pts_map = dict(self.cur_undistorted_pts_map)

    def published(self):
        """What the callers forward: only track_cnt > 1 (measurement_processor.cpp:212-223)."""
        out = {}
        for j in range(len(self.ids)):
            if self.track_cnt[j] > 1:
                out[self.ids[j]] = (float(self.cur_undistorted_pts[j][0]), float(self.cur_undistorted_pts[j][1]),
                                    1.0, float(self.cur_pts[j][0]), float(self.cur_pts[j][1]),
                                    float(self.pts_velocity[j][0]), float(self.pts_velocity[j][1]))
        return out


def edge_recovery(status, F, un_cur, un_next, next_pts, c: RefConfig) -> np.ndarray:
    """ft:236-285 — distance-aware recovery of rejected edge features; returns the updated status (uint8)."""
    status = np.array(status, np.uint8).reshape(-1)
    cx, cy = c.col / 2.0, c.row / 2.0
    ef = c.f_threshold_edge_factor
    if ef > 0.0 and F is not None and np.asarray(F).shape == (3, 3):
        r_max = np.sqrt(cx * cx + cy * cy)
        r_max_inv = 1.0 / r_max if r_max > 1e-6 else 0.0
        f = np.asarray(F, np.float64).ravel()
        for i in range(len(status)):
            if status[i]:
                continue
            dx, dy = float(next_pts[i][0]) - cx, float(next_pts[i][1]) - cy
            rr = np.sqrt(dx * dx + dy * dy) * r_max_inv
            if rr < 0.3:
                continue
            u1, v1 = float(un_cur[i][0]), float(un_cur[i][1])
            u2, v2 = float(un_next[i][0]), float(un_next[i][1])
            la = f[0] * u1 + f[1] * v1 + f[2]
            lb = f[3] * u1 + f[4] * v1 + f[5]
            lc = f[6] * u1 + f[7] * v1 + f[8]
            nab = np.sqrt(la * la + lb * lb)
            if nab < 1e-12:
                continue
            dist = abs(la * u2 + lb * v2 + lc) / nab
            if dist < c.f_threshold * (1.0 + ef * rr * rr):
                status[i] = 1
    return status


def config_from_synth(cfg: dict, **over) -> RefConfig:
    rc = RefConfig(row=cfg["height"], col=cfg["width"], focal_length=cfg["focal_length"],
                   max_cnt=cfg["max_cnt"], min_dist=cfg["min_dist"], fisheye=cfg["fisheye"],
                   lk_pyramid_levels=cfg["max_level"])
    for k, v in over.items():
        setattr(rc, k, v)
    return rc
