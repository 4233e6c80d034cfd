"""TEST INFRASTRUCTURE ONLY — the reference's ``frontend::FeatureManager``, restated statement by statement.

Follows ``/root/reference/src/frontend/feature_manager.cpp`` (quoted as ``fm:NNN``) with plain Python floats (IEEE
double, no fused multiply-add), in the reference's operation order, so that it is bit-equal to the compiled reference
(``oracle/_ref/libvsref_fm.so``, bound by ``oracle/ref_fm_binding.py``; pinned in tests/test_oracle_feature_manager.py).
The one library call of the path, ``Eigen::JacobiSVD`` (fm:182), is ``numpy.linalg.svd`` (LAPACK) on both sides.
Used as the checker of the device feature bank (``vs_bank_*``).  Never imported by the product.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

WINDOW_SIZE = 10  # include/utility/config.h:11


@dataclass
class FeaturePerId:  # include/frontend/feature_manager.h:49-68
    feature_id: int
    start_frame: int
    obs: list = field(default_factory=list)  # rows [x, y, z, u, v, vx, vy]  (FeaturePerFrame, feature_manager.h:25-47)
    used_num: int = 0
    estimated_depth: float = -1.0
    solve_flag: int = 0

    def end_frame(self) -> int:  # fm:5-7
        return self.start_frame + len(self.obs) - 1


def _mat_vec(m, v):  # small fixed product: every coefficient is the left-to-right sum of its products
    return [(m[i][0] * v[0] + m[i][1] * v[1]) + m[i][2] * v[2] for i in range(3)]


def _mat_mat(a, b):
    return [[(a[i][0] * b[0][j] + a[i][1] * b[1][j]) + a[i][2] * b[2][j] for j in range(3)] for i in range(3)]


def _transpose(m):
    return [[m[j][i] for j in range(3)] for i in range(3)]


def svd_v(a: np.ndarray) -> np.ndarray:
    """Right singular vectors of ``a`` (m x n, m >= n), singular values descending — JacobiSVD(ComputeThinV).matrixV()."""
    return np.linalg.svd(np.asarray(a, dtype=np.float64), full_matrices=False)[2].T


class FeatureManager:
    def __init__(self, min_parallax: float = 10.0, focal_length: float = 460.0, init_depth: float = 5.0):
        self.bank: list[FeaturePerId] = []  # feature_bank_ (a std::list, insertion order)
        self.last_track_num = 0
        self.min_parallax = min_parallax  # g_config.estimator.min_parallax
        self.focal_length = focal_length  # g_config.camera.focal_length
        self.init_depth = init_depth  # g_config.estimator.init_depth
        self.last_parallax = (0.0, 0)  # (parallax_sum, parallax_num) of the last call, for the tests

    def clear_state(self):  # fm:11-13
        self.bank = []

    def _countable(self, it: FeaturePerId) -> bool:
        it.used_num = len(it.obs)
        return it.used_num >= 2 and it.start_frame < WINDOW_SIZE - 2

    def get_feature_count(self) -> int:  # fm:15-25
        return sum(1 for it in self.bank if self._countable(it))

    def add_feature_and_check_parallax(self, frame_count: int, image: dict) -> bool:  # fm:27-73
        """``image``: {feature_id: 7 doubles}; iterated in ascending id order like the std::map behind common::ImageData."""
        parallax_sum = 0.0
        parallax_num = 0
        self.last_track_num = 0
        for feature_id in sorted(image):
            row = [float(v) for v in image[feature_id]]
            it = next((e for e in self.bank if e.feature_id == feature_id), None)  # find_if: the first match
            if it is None:
                self.bank.append(FeaturePerId(feature_id, frame_count, [row]))
            else:
                it.obs.append(row)
                self.last_track_num += 1
        self.last_parallax = (0.0, 0)
        if frame_count < 2 or self.last_track_num < 20:
            return True
        for it in self.bank:
            if it.start_frame <= frame_count - 2 and it.start_frame + len(it.obs) - 1 >= frame_count - 1:
                parallax_sum += self.compensated_parallax2(it, frame_count)
                parallax_num += 1
        self.last_parallax = (parallax_sum, parallax_num)
        if parallax_num == 0:
            return True
        return (parallax_sum / parallax_num) >= (self.min_parallax / self.focal_length)

    @staticmethod
    def compensated_parallax2(it: FeaturePerId, frame_count: int) -> float:  # fm:264-294
        fi = it.obs[frame_count - 2 - it.start_frame]
        fj = it.obs[frame_count - 1 - it.start_frame]
        u_j, v_j = fj[0], fj[1]
        dep_i = fi[2]
        u_i = fi[0] / dep_i
        v_i = fi[1] / dep_i
        du, dv = u_i - u_j, v_i - v_j
        # p_i_comp = p_i: the "compensated" terms repeat the same arithmetic
        du_comp, dv_comp = u_i - u_j, v_i - v_j
        return max(0.0, math.sqrt(min(du * du + dv * dv, du_comp * du_comp + dv_comp * dv_comp)))

    def get_corresponding(self, l: int, r: int) -> list:  # fm:75-92
        out = []
        for it in self.bank:
            if it.start_frame <= l and it.end_frame() >= r:
                out.append((it.obs[l - it.start_frame][:3], it.obs[r - it.start_frame][:3]))
        return out

    def set_depth(self, x):  # fm:94-110
        k = -1
        for it in self.bank:
            if not self._countable(it):
                continue
            k += 1
            depth = 1.0 / float(x[k])
            if depth < 0:
                it.solve_flag = 2
            else:
                it.estimated_depth = depth
                it.solve_flag = 1

    def remove_failures(self):  # fm:112-118
        self.bank = [it for it in self.bank if it.solve_flag != 2]

    def clear_depth(self, x):  # fm:120-128
        k = -1
        for it in self.bank:
            if not self._countable(it):
                continue
            k += 1
            it.estimated_depth = 1.0 / float(x[k])

    def get_depth_vector(self) -> list:  # fm:130-140
        return [1.0 / it.estimated_depth for it in self.bank if self._countable(it)]

    def triangulate_across_all_views(self, poses, t_ic, r_ic):  # fm:142-194
        """``poses``: WINDOW_SIZE + 1 entries (R 3x3 nested lists / array, P 3-vector) — sliding_window[i].R / .P."""
        R = [[[float(p[0][i][j]) for j in range(3)] for i in range(3)] for p in poses]
        P = [[float(p[1][i]) for i in range(3)] for p in poses]
        t_ic = [float(v) for v in t_ic]
        r_ic = [[float(r_ic[i][j]) for j in range(3)] for i in range(3)]
        for it in self.bank:
            if not self._countable(it):
                continue
            if it.estimated_depth > 0:
                continue
            imu_i = it.start_frame
            imu_j = imu_i - 1
            rt = _mat_vec(R[imu_i], t_ic)
            t0 = [P[imu_i][k] + rt[k] for k in range(3)]
            R0 = _mat_mat(R[imu_i], r_ic)
            R0t = _transpose(R0)
            A = np.zeros((2 * len(it.obs), 4))
            row = 0
            for ob in it.obs:
                imu_j += 1
                rt = _mat_vec(R[imu_j], t_ic)
                t1 = [P[imu_j][k] + rt[k] for k in range(3)]
                R1 = _mat_mat(R[imu_j], r_ic)
                rel_t = _mat_vec(R0t, [t1[k] - t0[k] for k in range(3)])
                rel_R = _mat_mat(R0t, R1)
                rel_Rt = _transpose(rel_R)
                neg = [[-rel_Rt[i][j] for j in range(3)] for i in range(3)]
                pt = _mat_vec(neg, rel_t)
                Pm = [rel_Rt[i] + [pt[i]] for i in range(3)]  # 3x4 [R^T | -R^T t]
                z = (ob[0] * ob[0] + ob[1] * ob[1]) + ob[2] * ob[2]
                f = ob[:3]
                if z > 0:
                    n = math.sqrt(z)
                    f = [ob[0] / n, ob[1] / n, ob[2] / n]
                for j in range(4):
                    A[row, j] = f[0] * Pm[2][j] - f[2] * Pm[0][j]
                    A[row + 1, j] = f[1] * Pm[2][j] - f[2] * Pm[1][j]
                row += 2
            v = svd_v(A)[:, 3]
            it.estimated_depth = float(v[2] / v[3])
            if it.estimated_depth < 0.1:
                it.estimated_depth = self.init_depth

    def remove_back_shift_depth(self, marg_R, marg_P, new_R, new_P):  # fm:208-235
        marg_R = [[float(marg_R[i][j]) for j in range(3)] for i in range(3)]
        new_Rt = _transpose([[float(new_R[i][j]) for j in range(3)] for i in range(3)])
        marg_P = [float(v) for v in marg_P]
        new_P = [float(v) for v in new_P]
        keep = []
        for it in self.bank:
            if it.start_frame != 0:
                it.start_frame -= 1
                keep.append(it)
                continue
            uv_i = it.obs[0][:3]
            del it.obs[0]
            if len(it.obs) < 2:
                continue
            pts_i = [uv_i[k] * it.estimated_depth for k in range(3)]
            rp = _mat_vec(marg_R, pts_i)
            w = [rp[k] + marg_P[k] for k in range(3)]
            pts_j = _mat_vec(new_Rt, [w[k] - new_P[k] for k in range(3)])
            dep_j = pts_j[2]
            it.estimated_depth = dep_j if dep_j > 0 else self.init_depth
            keep.append(it)
        self.bank = keep

    def remove_back(self):  # fm:237-249
        keep = []
        for it in self.bank:
            if it.start_frame != 0:
                it.start_frame -= 1
            else:
                del it.obs[0]
                if len(it.obs) == 0:
                    continue
            keep.append(it)
        self.bank = keep

    def remove_front(self, frame_count: int):  # fm:251-266
        keep = []
        for it in self.bank:
            if it.start_frame == frame_count:
                it.start_frame -= 1
            else:
                j = WINDOW_SIZE - 1 - it.start_frame
                if it.end_frame() < frame_count - 1:
                    keep.append(it)
                    continue
                del it.obs[j]
                if len(it.obs) == 0:
                    continue
            keep.append(it)
        self.bank = keep

    # ---- the bank as arrays (list order), the layout vs_bank_export uses ----
    def export(self, max_obs: int = WINDOW_SIZE + 1):
        n = len(self.bank)
        ids = np.array([it.feature_id for it in self.bank], dtype=np.int32)
        start = np.array([it.start_frame for it in self.bank], dtype=np.int32)
        n_obs = np.array([len(it.obs) for it in self.bank], dtype=np.int32)
        depth = np.array([it.estimated_depth for it in self.bank], dtype=np.float64)
        flag = np.array([it.solve_flag for it in self.bank], dtype=np.int32)
        obs = np.zeros((n, max_obs, 7))
        for e, it in enumerate(self.bank):
            for k, row in enumerate(it.obs[:max_obs]):
                obs[e, k] = row
        return ids, start, n_obs, depth, flag, obs
