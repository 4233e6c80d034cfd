"""TEST INFRASTRUCTURE ONLY — ctypes binding of oracle/_ref/libvsref_fm.so: the REFERENCE's own
``frontend::FeatureManager`` (``/root/reference/src/frontend/feature_manager.cpp`` compiled whole and unmodified; see
oracle/Makefile and oracle/ref_driver_fm.cpp).  ``CompiledRefFeatureManager`` has the surface of
``oracle.feature_manager.FeatureManager`` so a test can run both on the same frames.

What is NOT the reference here: Eigen is a stand-in (oracle/ref_shim/Eigen/Dense: fixed-size arithmetic in Eigen's
coefficient order) and ``Eigen::JacobiSVD`` is LAPACK (``numpy.linalg.svd``) through a callback.
Never imported by the product.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .feature_manager import WINDOW_SIZE, svd_v

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libvsref_fm.so")
_f64p = C.POINTER(C.c_double)
_i32p = C.POINTER(C.c_int)
_SVD_CB = C.CFUNCTYPE(None, C.c_int, C.c_int, _f64p, _f64p)
_lib = None
_keep = []


def available() -> bool:
    return os.path.exists(LIB_PATH)


def _svd_cb(m, n, a, v):
    A = np.ctypeslib.as_array(a, shape=(n, m)).T  # column-major in
    V = svd_v(A)
    out = np.ctypeslib.as_array(v, shape=(n, n))
    out[:, :] = V.T  # column-major out


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not available():
        raise RuntimeError(f"{LIB_PATH} missing: run `make -C oracle ref` where /root/reference exists")
    lib = C.CDLL(LIB_PATH)
    lib.vsref_fm_create.restype = C.c_void_p
    for name in ("vsref_fm_destroy", "vsref_fm_clear", "vsref_fm_remove_back", "vsref_fm_remove_failures"):
        getattr(lib, name).argtypes = [C.c_void_p]
        getattr(lib, name).restype = None
    lib.vsref_fm_set_config.argtypes = [C.c_double] * 3
    lib.vsref_fm_set_config.restype = None
    lib.vsref_fm_add_frame.argtypes = [C.c_void_p, C.c_int, C.c_int, _i32p, _f64p, _i32p]
    lib.vsref_fm_bank_size.argtypes = [C.c_void_p]
    lib.vsref_fm_feature_count.argtypes = [C.c_void_p]
    lib.vsref_fm_export.argtypes = [C.c_void_p, C.c_int, _i32p, _i32p, _i32p, _f64p, _i32p, _f64p]
    lib.vsref_fm_export.restype = None
    lib.vsref_fm_remove_front.argtypes = [C.c_void_p, C.c_int]
    lib.vsref_fm_remove_front.restype = None
    lib.vsref_fm_remove_back_shift_depth.argtypes = [C.c_void_p, _f64p, _f64p, _f64p, _f64p]
    lib.vsref_fm_remove_back_shift_depth.restype = None
    lib.vsref_fm_triangulate.argtypes = [C.c_void_p, _f64p, _f64p, _f64p]
    lib.vsref_fm_triangulate.restype = None
    lib.vsref_fm_get_depth_vector.argtypes = [C.c_void_p, _f64p]
    lib.vsref_fm_set_depth.argtypes = [C.c_void_p, _f64p, C.c_int]
    lib.vsref_fm_set_depth.restype = None
    lib.vsref_fm_clear_depth.argtypes = [C.c_void_p, _f64p, C.c_int]
    lib.vsref_fm_clear_depth.restype = None
    lib.vsref_fm_get_corresponding.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, _f64p]
    cb = _SVD_CB(_svd_cb)
    _keep.append(cb)
    lib.vsref_fm_set_svd_callback.argtypes = [_SVD_CB]
    lib.vsref_fm_set_svd_callback.restype = None
    lib.vsref_fm_set_svd_callback(cb)
    assert lib.vsref_fm_window_size() == WINDOW_SIZE
    _lib = lib
    return lib


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(_f64p)


class CompiledRefFeatureManager:
    """One compiled ``frontend::FeatureManager``.  ``utility::g_config`` is process-wide like in the reference: the three
    values it reads are written before every call that reads them."""

    def __init__(self, min_parallax: float = 10.0, focal_length: float = 460.0, init_depth: float = 5.0):
        self.lib = load()
        self.h = C.c_void_p(self.lib.vsref_fm_create())
        self.cfg = (min_parallax, focal_length, init_depth)
        self.last_track_num = 0

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.vsref_fm_destroy(self.h)
            self.h = None

    def _cfg(self):
        self.lib.vsref_fm_set_config(*self.cfg)

    def clear_state(self):
        self.lib.vsref_fm_clear(self.h)

    def add_feature_and_check_parallax(self, frame_count: int, image: dict) -> bool:
        self._cfg()
        ids = np.array(list(image.keys()), dtype=np.int32)
        feat, fp = _d(np.array([image[k] for k in image], dtype=np.float64).reshape(-1, 7))
        ltn = C.c_int(0)
        key = self.lib.vsref_fm_add_frame(self.h, frame_count, len(ids), ids.ctypes.data_as(_i32p), fp, C.byref(ltn))
        self.last_track_num = ltn.value
        return bool(key)

    def get_feature_count(self) -> int:
        return self.lib.vsref_fm_feature_count(self.h)

    def export(self, max_obs: int = WINDOW_SIZE + 1):
        n = self.lib.vsref_fm_bank_size(self.h)
        ids = np.zeros(n, np.int32); start = np.zeros(n, np.int32); n_obs = np.zeros(n, np.int32)
        depth = np.zeros(n); flag = np.zeros(n, np.int32); obs = np.zeros((n, max_obs, 7))
        self.lib.vsref_fm_export(self.h, max_obs, ids.ctypes.data_as(_i32p), start.ctypes.data_as(_i32p),
                                 n_obs.ctypes.data_as(_i32p), depth.ctypes.data_as(_f64p), flag.ctypes.data_as(_i32p),
                                 obs.ctypes.data_as(_f64p))
        return ids, start, n_obs, depth, flag, obs

    def remove_back(self):
        self.lib.vsref_fm_remove_back(self.h)

    def remove_front(self, frame_count: int):
        self.lib.vsref_fm_remove_front(self.h, frame_count)

    def remove_failures(self):
        self.lib.vsref_fm_remove_failures(self.h)

    def remove_back_shift_depth(self, marg_R, marg_P, new_R, new_P):
        self._cfg()
        a, ap = _d(marg_R); b, bp = _d(marg_P); c, cp = _d(new_R); d, dp = _d(new_P)
        self.lib.vsref_fm_remove_back_shift_depth(self.h, ap, bp, cp, dp)

    def triangulate_across_all_views(self, poses, t_ic, r_ic):
        self._cfg()
        flat = np.zeros((WINDOW_SIZE + 1, 12))
        for i, (R, P) in enumerate(poses):
            flat[i, :9] = np.asarray(R, dtype=np.float64).reshape(9)
            flat[i, 9:] = np.asarray(P, dtype=np.float64)
        f, fp = _d(flat); t, tp = _d(t_ic); r, rp = _d(r_ic)
        self.lib.vsref_fm_triangulate(self.h, fp, tp, rp)

    def get_depth_vector(self) -> list:
        out = np.zeros(max(self.lib.vsref_fm_bank_size(self.h), 1))
        n = self.lib.vsref_fm_get_depth_vector(self.h, out.ctypes.data_as(_f64p))
        return list(out[:n])

    def set_depth(self, x):
        a, ap = _d(x)
        self.lib.vsref_fm_set_depth(self.h, ap, len(a))

    def clear_depth(self, x):
        a, ap = _d(x)
        self.lib.vsref_fm_clear_depth(self.h, ap, len(a))

    def get_corresponding(self, l: int, r: int) -> list:
        cap = max(self.lib.vsref_fm_bank_size(self.h), 1)
        out = np.zeros((cap, 6))
        n = self.lib.vsref_fm_get_corresponding(self.h, l, r, cap, out.ctypes.data_as(_f64p))
        return [(list(out[i, :3]), list(out[i, 3:])) for i in range(n)]
