"""TEST INFRASTRUCTURE ONLY — a synthetic estimator loop around a FeatureManager, the way
``/root/reference/src/backend/estimator.cpp`` drives it (``processImage`` :136-150, ``solveOdometry`` :342-350,
``slideWindowOldKeyframe`` :313-340, ``slideWindowNewGeneralFrame`` :296-311, ``removeFailures`` :206,254).

A camera moves through a cloud of 3-D points; every frame publishes ``{id: [x, y, 1, u, v, vx, vy]}`` for the visible
points (tracks break at random and come back under a new id, like the tracker's), the manager decides key frame or not,
and once the window is full the loop triangulates, sometimes feeds the depth vector back with a few negative entries
(``setDepth`` + ``removeFailures``) and slides the window.  The same ``steps()`` generator feeds the Python restatement,
the compiled reference and the device feature bank, so their states can be compared after every operation.
"""
from __future__ import annotations

import math

import numpy as np

from .feature_manager import WINDOW_SIZE


def _rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = math.cos(rx), math.sin(rx), math.cos(ry), math.sin(ry), math.cos(rz), math.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


class Scenario:
    """Deterministic per seed.  ``speed`` scales the camera translation per frame (small speed => low parallax => frames
    that are not key frames)."""

    def __init__(self, seed: int, n_points: int = 260, max_feat: int = 150, speed: float = 0.06, drop: float = 0.03,
                 focal: float = 460.0, dt: float = 0.05):
        self.rng = np.random.default_rng(seed)
        self.pts = np.column_stack([self.rng.uniform(-6, 6, n_points), self.rng.uniform(-4, 4, n_points),
                                    self.rng.uniform(3, 12, n_points)])
        self.track_id = -np.ones(n_points, dtype=np.int64)
        self.prev_uv = {}
        self.next_id = 0
        self.max_feat, self.speed, self.drop, self.focal, self.dt = max_feat, speed, drop, focal, dt
        self.t_ic = np.array([0.02, -0.01, 0.005])
        self.r_ic = _rot(0.01, -0.02, 0.015)
        self.k = 0

    def pose(self, k: int):
        """IMU pose of frame k (R world<-imu, P)."""
        s = self.speed
        P = np.array([s * k, 0.3 * s * math.sin(0.3 * k), 0.1 * s * k])
        R = _rot(0.004 * k, 0.01 * math.sin(0.2 * k), 0.003 * k)
        return R, P

    def frame(self):
        """Next frame's published records {id: 7 doubles} and its pose."""
        k = self.k
        self.k += 1
        R, P = self.pose(k)
        Rc = R @ self.r_ic
        tc = P + R @ self.t_ic
        pc = (self.pts - tc) @ Rc  # camera coordinates: Rc^T (X - tc)
        vis = (pc[:, 2] > 0.5) & (np.abs(pc[:, 0] / pc[:, 2]) < 0.8) & (np.abs(pc[:, 1] / pc[:, 2]) < 0.5)
        lost = self.rng.random(len(self.pts)) < self.drop
        self.track_id[~vis | lost] = -1
        image = {}
        n_live = 0
        for i in np.nonzero(vis & ~lost)[0]:
            if n_live >= self.max_feat:
                self.track_id[i] = -1
                continue
            n_live += 1
            fresh = self.track_id[i] < 0
            if fresh:
                self.track_id[i] = self.next_id
                self.next_id += 1
            fid = int(self.track_id[i])
            # the tracker publishes float32 undistorted points, pixel coordinates and velocities
            x = float(np.float32(pc[i, 0] / pc[i, 2] + self.rng.normal(0, 2e-4)))
            y = float(np.float32(pc[i, 1] / pc[i, 2] + self.rng.normal(0, 2e-4)))
            u = float(np.float32(x * self.focal + 376.0))
            v = float(np.float32(y * self.focal + 240.0))
            if fid in self.prev_uv and not fresh:
                vx = float(np.float32((x - self.prev_uv[fid][0]) / self.dt))
                vy = float(np.float32((y - self.prev_uv[fid][1]) / self.dt))
            else:
                vx = vy = 0.0
            self.prev_uv[fid] = (x, y)
            image[fid] = [x, y, 1.0, u, v, vx, vy]
        return image, (R, P)


def steps(seed: int, n_frames: int, **kw):
    """Yield the estimator's operations: ("add", frame_count, image) | ("triangulate", poses, t_ic, r_ic) |
    ("set_depth", x_fn) + ("remove_failures",) | ("remove_back",) | ("remove_back_shift_depth", R0, P0, R1, P1) |
    ("remove_front", frame_count).  After "add" the caller sends back the key-frame decision with ``gen.send(key)``;
    after "set_depth" prep ("depth_vector") it sends the current depth vector."""
    sc = Scenario(seed, **kw)
    rng = np.random.default_rng(seed + 7919)
    frame_count = 0
    window = [None] * (WINDOW_SIZE + 1)
    nonlinear = False
    for _ in range(n_frames):
        image, pose = sc.frame()
        window[frame_count] = pose
        key = yield ("add", frame_count, image)
        if frame_count < WINDOW_SIZE:
            frame_count += 1
            continue
        poses = [w if w is not None else window[frame_count] for w in window]
        if nonlinear:
            yield ("triangulate", poses, sc.t_ic, sc.r_ic)
            if rng.random() < 0.4:  # the optimiser's result comes back: a few inverse depths turn negative
                dep = yield ("depth_vector",)
                x = np.array(dep, dtype=np.float64) * (1.0 + rng.normal(0, 0.02, len(dep)))
                neg = rng.random(len(dep)) < 0.05
                x[neg] = -np.abs(x[neg])
                yield ("set_depth", x)
                yield ("remove_failures",)
        if key:
            back_R0, back_P0 = window[0]
            window = window[1:] + [window[WINDOW_SIZE]]
            if nonlinear:
                R0 = back_R0 @ sc.r_ic
                R1 = window[0][0] @ sc.r_ic
                P0 = back_P0 + back_R0 @ sc.t_ic
                P1 = window[0][1] + window[0][0] @ sc.t_ic
                yield ("remove_back_shift_depth", R0, P0, R1, P1)
            else:
                yield ("remove_back",)
        else:
            window[WINDOW_SIZE - 1] = window[WINDOW_SIZE]
            yield ("remove_front", frame_count)
        nonlinear = True  # the first full window initialises the estimator (solver_flag_ = NON_LINEAR)


def drive(fm, seed: int, n_frames: int, on_step=None, **kw):
    """Run the scenario on one manager (any object with the FeatureManager surface).  ``on_step(op, fm, result)`` is called
    after every operation.  Returns the list of key-frame decisions."""
    gen = steps(seed, n_frames, **kw)
    keys = []
    op = next(gen, None)
    while op is not None:
        result = None
        send = None
        if op[0] == "add":
            result = fm.add_feature_and_check_parallax(op[1], op[2])
            keys.append(result)
            send = result
        elif op[0] == "triangulate":
            fm.triangulate_across_all_views(op[1], op[2], op[3])
        elif op[0] == "depth_vector":
            result = fm.get_depth_vector()
            send = result
        elif op[0] == "set_depth":
            fm.set_depth(op[1])
        elif op[0] == "remove_failures":
            fm.remove_failures()
        elif op[0] == "remove_back":
            fm.remove_back()
        elif op[0] == "remove_back_shift_depth":
            fm.remove_back_shift_depth(op[1], op[2], op[3], op[4])
        elif op[0] == "remove_front":
            fm.remove_front(op[1])
        if on_step:
            on_step(op, fm, result)
        try:
            op = gen.send(send)
        except StopIteration:
            op = None
    return keys
