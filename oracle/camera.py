"""TEST INFRASTRUCTURE ONLY — float64 restatement of ``Camera::liftProjective``.

Follows, for each model:
  PINHOLE         /root/reference/src/common/camera_models/PinholeCamera.cc:358-410 (+distortion :532-547,
                  ctor :210-223)
  KANNALA_BRANDT  EquidistantCamera.cc:337-349 (+backprojectSymmetric :574-652: smallest non-negative
                  real eigenvalue of the companion matrix)
  MEI             CataCamera.cc:444-503 (+distortion :629-644, ctor :236-249)
  SCARAMUZZA      ScaramuzzaCamera.cc:575-594 (note: the ray's x,y are the *un-affined* centred pixel)

Model enum values follow include/common/camera_models/Camera.h:15
(KANNALA_BRANDT=0, MEI=1, PINHOLE=2, SCARAMUZZA=3).
"""
from __future__ import annotations

import math

import numpy as np

KANNALA_BRANDT, MEI, PINHOLE, SCARAMUZZA = 0, 1, 2, 3


class Camera:
    def __init__(self, model: int, fx=1.0, fy=1.0, cx=0.0, cy=0.0, d=(0.0, 0.0, 0.0, 0.0), xi=0.0,
                 poly=(0.0,) * 5, affine=(1.0, 0.0, 0.0)):
        self.model = int(model)
        self.fx, self.fy, self.cx, self.cy = float(fx), float(fy), float(cx), float(cy)
        self.d = tuple(float(v) for v in d)   # pinhole/mei: k1,k2,p1,p2 ; KB: k2,k3,k4,k5
        self.xi = float(xi)
        self.poly = tuple(float(v) for v in poly)
        self.C, self.D, self.E = (float(v) for v in affine)
        if self.model != SCARAMUZZA:
            self.inv_K11 = 1.0 / self.fx
            self.inv_K13 = -self.cx / self.fx
            self.inv_K22 = 1.0 / self.fy
            self.inv_K23 = -self.cy / self.fy
        else:
            self.inv_scale = 1.0 / (self.C - self.D * self.E)

    # --- radtan helper (PinholeCamera.cc:532-547 / CataCamera.cc:629-644) ------------------------
    def _distortion(self, ux, uy):
        k1, k2, p1, p2 = self.d
        mx2, my2, mxy = ux * ux, uy * uy, ux * uy
        rho2 = mx2 + my2
        rad = k1 * rho2 + k2 * rho2 * rho2
        return (ux * rad + 2.0 * p1 * mxy + p2 * (rho2 + 2.0 * mx2),
                uy * rad + 2.0 * p2 * mxy + p1 * (rho2 + 2.0 * my2))

    def _undistort_radtan(self, mx_d, my_d):
        if all(v == 0.0 for v in self.d):
            return mx_d, my_d
        dx, dy = self._distortion(mx_d, my_d)
        mx_u, my_u = mx_d - dx, my_d - dy
        for _ in range(1, 8):
            dx, dy = self._distortion(mx_u, my_u)
            mx_u, my_u = mx_d - dx, my_d - dy
        return mx_u, my_u

    def _kb_theta(self, r):
        k2, k3, k4, k5 = self.d
        npow = 9
        if k5 == 0.0:
            npow -= 2
        if k4 == 0.0:
            npow -= 2
        if k3 == 0.0:
            npow -= 2
        if k2 == 0.0:
            npow -= 2
        if npow == 1:
            return r
        coeffs = np.zeros(npow + 1)
        coeffs[0] = -r
        coeffs[1] = 1.0
        if npow >= 3:
            coeffs[3] = k2
        if npow >= 5:
            coeffs[5] = k3
        if npow >= 7:
            coeffs[7] = k4
        if npow >= 9:
            coeffs[9] = k5
        A = np.zeros((npow, npow))
        A[1:, :npow - 1] = np.eye(npow - 1)
        A[:, npow - 1] = -coeffs[:npow] / coeffs[npow]
        ev = np.linalg.eigvals(A)
        tol = 1e-10
        thetas = []
        for e in ev:
            if abs(e.imag) > tol:
                continue
            t = e.real
            if t < -tol:
                continue
            if t < 0.0:
                t = 0.0
            thetas.append(t)
        return min(thetas) if thetas else r

    def lift_projective(self, u: float, v: float):
        """pixel -> projective ray (float64 triple)."""
        if self.model == SCARAMUZZA:
            xc0, xc1 = u - self.cx, v - self.cy
            xa0 = self.inv_scale * (xc0 - self.D * xc1)
            xa1 = self.inv_scale * (-self.E * xc0 + self.C * xc1)
            phi = math.sqrt(xa0 * xa0 + xa1 * xa1)
            phi_i, z = 1.0, 0.0
            for c in self.poly:
                z += phi_i * c
                phi_i *= phi
            return xc0, xc1, -z
        mx = self.inv_K11 * u + self.inv_K13
        my = self.inv_K22 * v + self.inv_K23
        if self.model == PINHOLE:
            mx_u, my_u = self._undistort_radtan(mx, my)
            return mx_u, my_u, 1.0
        if self.model == MEI:
            mx_u, my_u = self._undistort_radtan(mx, my)
            xi = self.xi
            if xi == 1.0:
                return mx_u, my_u, (1.0 - mx_u * mx_u - my_u * my_u) / 2.0
            rho2 = mx_u * mx_u + my_u * my_u
            disc = 1.0 + (1.0 - xi * xi) * rho2       # negative outside the mirror's field of view: std::sqrt gives NaN
            root = math.sqrt(disc) if disc >= 0.0 else float("nan")
            return mx_u, my_u, 1.0 - xi * (rho2 + 1.0) / (xi + root)
        # KANNALA_BRANDT
        r = math.sqrt(mx * mx + my * my)
        phi = 0.0 if r < 1e-10 else math.atan2(my, mx)
        theta = self._kb_theta(r)
        return (math.sin(theta) * math.cos(phi), math.sin(theta) * math.sin(phi), math.cos(theta))

    def lift_batch(self, pts) -> np.ndarray:
        """Vectorised ``lift_projective`` over (N,2) pixels -> (N,3) float64 (same arithmetic order)."""
        pts = np.asarray(pts, np.float64).reshape(-1, 2)
        n = len(pts)
        if n == 0:
            return np.zeros((0, 3))
        if self.model != KANNALA_BRANDT:
            return np.array([self.lift_projective(float(u), float(v)) for u, v in pts])
        mx = self.inv_K11 * pts[:, 0] + self.inv_K13
        my = self.inv_K22 * pts[:, 1] + self.inv_K23
        r = np.sqrt(mx * mx + my * my)
        phi = np.where(r < 1e-10, 0.0, np.arctan2(my, mx))
        k2, k3, k4, k5 = self.d
        npow = 9 - 2 * sum(1 for k in (k5, k4, k3, k2) if k == 0.0)
        if npow == 1:
            theta = r
        else:
            coeffs = np.zeros((n, npow + 1))
            coeffs[:, 0] = -r
            coeffs[:, 1] = 1.0
            for pw, k in ((3, k2), (5, k3), (7, k4), (9, k5)):
                if npow >= pw:
                    coeffs[:, pw] = k
            A = np.zeros((n, npow, npow))
            A[:, 1:, :npow - 1] = np.eye(npow - 1)
            A[:, :, npow - 1] = -coeffs[:, :npow] / coeffs[:, npow:npow + 1]
            ev = np.linalg.eigvals(A)
            ok = (np.abs(ev.imag) <= 1e-10) & (ev.real >= -1e-10)
            t = np.where(ok, np.maximum(ev.real, 0.0), np.inf)
            theta = t.min(axis=1)
            theta = np.where(np.isfinite(theta), theta, r)
        return np.stack([np.sin(theta) * np.cos(phi), np.sin(theta) * np.sin(phi), np.cos(theta)], 1)

    def undistort_f32(self, pts) -> np.ndarray:
        """(x/z, y/z) cast to float32 — what ``undistortedPoints`` stores (feature_tracker.cpp:343-347)."""
        P = self.lift_batch(pts)
        if len(P) == 0:
            return np.zeros((0, 2), np.float32)
        return np.stack([P[:, 0] / P[:, 2], P[:, 1] / P[:, 2]], 1).astype(np.float32)


def from_config(cfg: dict) -> Camera:
    """Camera from a ``mobile_slam_b200.synth.CONFIGS`` entry."""
    model = {"PINHOLE": PINHOLE, "KANNALA_BRANDT": KANNALA_BRANDT, "MEI": MEI,
             "SCARAMUZZA": SCARAMUZZA}[cfg["model"]]
    fx, fy, cx, cy = cfg["intr"]
    return Camera(model, fx, fy, cx, cy, cfg["dist"], xi=cfg.get("xi", 0.0),
                  poly=cfg.get("poly", (0.0,) * 5), affine=cfg.get("affine", (1.0, 0.0, 0.0)))
