"""TEST INFRASTRUCTURE ONLY — ctypes binding of oracle/_ref/libvsref.so: the REFERENCE's own
``frontend::FeatureTracker`` (``/root/reference/src/frontend/feature_tracker.cpp`` compiled whole and unmodified, with
the ``liftProjective`` functions of the reference's camera models), its OpenCV calls forwarded to the real cv2 wheel.

``CompiledRefTracker`` has the surface of ``oracle.cv_reference.RefFeatureTracker`` (same constructor arguments, same
public vectors), so a test can run both and the CUDA tracker on the same frames:

    reference C++ glue + reference camera code + cv2 arithmetic   (this file)       <- pins
    Python restatement of the glue + camera + cv2 arithmetic      (cv_reference.py)  <- checks
    CUDA kernels                                                  (the product)

What is NOT the reference here, and why: OpenCV itself is the cv2 4.13.0 wheel (the reference builds against an
unpinned "OpenCV 4.x"); ``Eigen::EigenSolver`` in the Kannala-Brandt back-projection is LAPACK (numpy) behind a
stand-in class (Eigen is absent); ``cv::FileStorage`` (camera YAML) is cv2.FileStorage.  See oracle/ref_shim/ and
oracle/Makefile.

Never imported by the product; bench.py may use it only in the CPU-baseline / ``--impl reference`` leg.
"""
from __future__ import annotations

import ctypes as C
import os

import cv2
import numpy as np

from . import camera as cam
from .cv_reference import RefConfig

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libvsref.so")

_u8p = C.POINTER(C.c_ubyte)
_f32p = C.POINTER(C.c_float)
_f64p = C.POINTER(C.c_double)
_i32p = C.POINTER(C.c_int)


class _Callbacks(C.Structure):
    _fields_ = [
        ("clahe_apply", C.CFUNCTYPE(None, C.c_double, C.c_int, C.c_int, _u8p, C.c_int, C.c_int, C.c_long, _u8p)),
        ("lk", C.CFUNCTYPE(None, _u8p, C.c_long, _u8p, C.c_long, C.c_int, C.c_int, _f32p, C.c_int, _f32p, _u8p, _f32p,
                           C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double)),
        ("gftt", C.CFUNCTYPE(C.c_int, _u8p, C.c_long, _u8p, C.c_long, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double,
                             _f32p)),
        ("find_f", C.CFUNCTYPE(C.c_int, _f32p, _f32p, C.c_int, C.c_int, C.c_double, C.c_double, _f64p, _u8p, _i32p)),
        ("circle", C.CFUNCTYPE(None, _u8p, C.c_long, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int)),
        ("imread_gray", C.CFUNCTYPE(None, C.c_char_p, _u8p, _i32p, _i32p)),
        ("yaml_get", C.CFUNCTYPE(C.c_int, C.c_char_p, C.c_char_p, C.c_char_p, _f64p, C.POINTER(C.c_char), C.c_int)),
        ("eigvals", C.CFUNCTYPE(None, C.c_int, _f64p, _f64p, _f64p)),
    ]


class _Config(C.Structure):
    _fields_ = [("row", C.c_double), ("col", C.c_double), ("focal_length", C.c_double),
                ("max_cnt", C.c_int), ("min_dist", C.c_int), ("f_threshold", C.c_double),
                ("equalize", C.c_int), ("fisheye", C.c_int),
                ("lk_window_size", C.c_int), ("lk_pyramid_levels", C.c_int), ("lk_iterations", C.c_int),
                ("lk_eps", C.c_double), ("f_threshold_edge_factor", C.c_double)]


def _img(ptr, rows, cols, step):
    """numpy view (rows, cols) of a strided u8 image owned by the C++ side."""
    flat = np.ctypeslib.as_array(ptr, shape=((rows - 1) * step + cols,))
    return np.lib.stride_tricks.as_strided(flat, shape=(rows, cols), strides=(step, 1))


_lib = None
_cb_keepalive = None
_active = None          # the tracker whose C++ call is running (callbacks record stage dumps into it)
_cb_error = None
_clahe_cache = {}


def available() -> bool:
    return os.path.exists(LIB_PATH)


def _guard(fn):
    def wrapped(*a):
        global _cb_error
        try:
            return fn(*a)
        except BaseException as e:  # an exception must not unwind through the C++ frames
            if _cb_error is None:
                _cb_error = e
            return 0
    return wrapped


@_guard
def _cb_clahe(clip, tx, ty, src, rows, cols, step, dst):
    key = (clip, tx, ty)
    if key not in _clahe_cache:
        _clahe_cache[key] = cv2.createCLAHE(clip, (tx, ty))
    out = _clahe_cache[key].apply(np.ascontiguousarray(_img(src, rows, cols, step)))
    np.ctypeslib.as_array(dst, shape=(rows, cols))[:] = out
    if _active is not None:
        _active.stage["clahe"] = out


@_guard
def _cb_lk(prev, pstep, nxt, nstep, rows, cols, ppts, n, npts, status, err, ww, wh, max_level, ctype, ccount, ceps):
    a = np.ascontiguousarray(_img(prev, rows, cols, pstep))
    b = np.ascontiguousarray(_img(nxt, rows, cols, nstep))
    p = np.ctypeslib.as_array(ppts, shape=(n, 2)).copy()
    q, st, er = cv2.calcOpticalFlowPyrLK(a, b, p.reshape(-1, 1, 2), None, winSize=(ww, wh), maxLevel=max_level,
                                         criteria=(ctype, ccount, ceps))
    np.ctypeslib.as_array(npts, shape=(n, 2))[:] = q.reshape(-1, 2)
    np.ctypeslib.as_array(status, shape=(n,))[:] = st.reshape(-1)
    np.ctypeslib.as_array(err, shape=(n,))[:] = er.reshape(-1)
    if _active is not None:
        _active.stage.update(lk_in=p, lk_out=q.reshape(-1, 2).copy(), lk_status_raw=st.reshape(-1).astype(np.uint8))


@_guard
def _cb_gftt(img, step, mask, mstep, rows, cols, max_corners, quality, min_dist, out):
    im = np.ascontiguousarray(_img(img, rows, cols, step))
    mk = np.ascontiguousarray(_img(mask, rows, cols, mstep)) if mask else None
    r = cv2.goodFeaturesToTrack(im, max_corners, quality, min_dist, mask=mk)
    pts = np.zeros((0, 2), np.float32) if r is None else r.reshape(-1, 2)
    np.ctypeslib.as_array(out, shape=(max(max_corners, 1), 2))[:len(pts)] = pts
    if _active is not None:
        _active.stage.update(n_pts=pts.copy(), mask=None if mk is None else mk.copy())
    return len(pts)


@_guard
def _cb_find_f(p1, p2, n, method, thr, conf, f9, status, has_status):
    a = np.ctypeslib.as_array(p1, shape=(n, 2)).copy()
    b = np.ctypeslib.as_array(p2, shape=(n, 2)).copy()
    st_out = np.ctypeslib.as_array(status, shape=(n,))
    t = _active
    if t is not None:
        t.stage["ransac_ran"] = True
        t.stage["ransac_pts"] = (a, b)
    if t is not None and not t.cfg.use_ransac:
        # stage isolation (RefConfig.use_ransac = False): keep every point, as if rejectWithFundamentalMatrix were skipped
        st_out[:] = 1
        has_status[0] = 1
        return 0
    fn = t.find_f if (t is not None and t.find_f is not None) else (
        lambda u, v: cv2.findFundamentalMat(u, v, method, thr, conf))
    F, st = fn(a, b)
    if st is not None:
        st_out[:] = np.asarray(st).reshape(-1)
        has_status[0] = 1
    else:
        has_status[0] = 0
    if t is not None:
        t.stage["F"] = None if F is None else np.array(F, np.float64)
    if F is None or np.asarray(F).shape != (3, 3):
        return 0
    np.ctypeslib.as_array(f9, shape=(9,))[:] = np.asarray(F, np.float64).ravel()
    return 1


@_guard
def _cb_circle(img, step, rows, cols, cx, cy, radius, color, thickness):
    # cv2.circle needs a contiguous array: draw on a copy of the touched rows and write them back
    view = _img(img, rows, cols, step)
    if step == cols:
        cv2.circle(np.ctypeslib.as_array(img, shape=(rows, cols)), (cx, cy), radius, color, thickness)
    else:
        tmp = np.ascontiguousarray(view)
        cv2.circle(tmp, (cx, cy), radius, color, thickness)
        view[:] = tmp


@_guard
def _cb_imread(path, out, rows, cols):
    im = cv2.imread(path.decode(), cv2.IMREAD_GRAYSCALE)
    if im is None:
        rows[0] = cols[0] = 0
        return
    rows[0], cols[0] = im.shape
    if out:
        np.ctypeslib.as_array(out, shape=im.shape)[:] = im


_yaml_cache = {}


@_guard
def _cb_yaml_get(path, k1, k2, num, sbuf, cap):
    """cv::FileStorage(path, READ)[k1][k2] answered by the real cv2.FileStorage"""
    path, k1, k2 = path.decode(), k1.decode(), k2.decode()
    fs = _yaml_cache.get(path)
    if fs is None:
        try:
            fs = cv2.FileStorage(path, cv2.FILE_STORAGE_READ)
        except cv2.error:
            return -1
        if not fs.isOpened():
            return -1
        _yaml_cache[path] = fs
    if not k1:
        return 0
    node = fs.getNode(k1)
    if k2:
        if node.empty() or not node.isMap():
            return 0
        node = node.getNode(k2)
    if node.empty() or node.isNone():
        return 0
    if node.isInt():
        num[0] = float(int(node.real()))
        return 1
    if node.isReal():
        num[0] = float(node.real())
        return 2
    if node.isString():
        b = node.string().encode()[: cap - 1]
        C.memmove(sbuf, b + b"\0", len(b) + 1)
        return 3
    return 0


@_guard
def _cb_eigvals(n, a, wr, wi):
    A = np.ctypeslib.as_array(a, shape=(n, n)).T.copy()     # column-major in, numpy is row-major
    ev = np.linalg.eigvals(A)
    np.ctypeslib.as_array(wr, shape=(n,))[:] = ev.real
    np.ctypeslib.as_array(wi, shape=(n,))[:] = ev.imag


def load():
    global _lib, _cb_keepalive
    if _lib is not None:
        return _lib
    if not available():
        raise FileNotFoundError(f"{LIB_PATH} is missing: run `make -C oracle ref` where /root/reference exists")
    lib = C.CDLL(LIB_PATH)
    lib.vsref_create.restype = C.c_void_p
    lib.vsref_destroy.argtypes = [C.c_void_p]
    lib.vsref_set_callbacks.argtypes = [C.POINTER(_Callbacks)]
    lib.vsref_set_config.argtypes = [C.POINTER(_Config), C.c_char_p]
    lib.vsref_set_n_id.argtypes = [C.c_int]
    lib.vsref_get_n_id.restype = C.c_int
    lib.vsref_set_intrinsic.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_double] * 8
    lib.vsref_set_camera_mei.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_double] * 9
    lib.vsref_set_camera_ocam.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_double] * 5 + [_f64p]
    lib.vsref_detect_and_track.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_long, C.c_double]
    lib.vsref_update_ids.argtypes = [C.c_void_p]
    lib.vsref_update_ids.restype = C.c_int
    lib.vsref_count.argtypes = [C.c_void_p]
    lib.vsref_count.restype = C.c_int
    lib.vsref_n_new.argtypes = [C.c_void_p]
    lib.vsref_n_new.restype = C.c_int
    lib.vsref_get.argtypes = [C.c_void_p] + [C.c_void_p] * 5
    lib.vsref_get_mask.argtypes = [C.c_void_p, C.c_void_p]
    lib.vsref_get_mask.restype = C.c_int
    lib.vsref_lift_projective.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.vsref_read_intrinsic.argtypes = [C.c_void_p, C.c_char_p]
    lib.vsref_read_intrinsic.restype = C.c_int
    lib.vsref_camera_parameters.argtypes = [C.c_void_p, C.c_void_p, C.c_int, _i32p, _i32p]
    lib.vsref_camera_parameters.restype = C.c_int
    lib.vsref_in_border.argtypes = [C.c_float, C.c_float]
    lib.vsref_in_border.restype = C.c_int
    cbs = _Callbacks()
    fns = dict(clahe_apply=_cb_clahe, lk=_cb_lk, gftt=_cb_gftt, find_f=_cb_find_f, circle=_cb_circle,
               imread_gray=_cb_imread, yaml_get=_cb_yaml_get, eigvals=_cb_eigvals)
    for name, ctype in _Callbacks._fields_:
        setattr(cbs, name, ctype(fns[name]))
    _cb_keepalive = cbs
    lib.vsref_set_callbacks(C.byref(cbs))
    _lib = lib
    return lib


def read_intrinsic_parameters(yaml_path: str):
    """``FeatureTracker::readIntrinsicParameter`` of the compiled reference (CameraFactory::generateCameraFromYamlFile +
    the model's Parameters::readFromYamlFile, cv::FileStorage served by cv2.FileStorage).  Returns
    ``(model_type, width, height, parameter vector as Camera::writeParameters orders it)`` or ``None`` when the factory
    returns a null camera (unknown model / unreadable file)."""
    lib = load()
    global _active, _cb_error
    _active, _cb_error = None, None
    _yaml_cache.pop(yaml_path, None)
    h = lib.vsref_create()
    try:
        mt = lib.vsref_read_intrinsic(h, yaml_path.encode())
        if _cb_error is not None:
            e, _cb_error = _cb_error, None
            raise e
        if mt < 0:
            return None
        buf = np.zeros(64)
        w, hh = C.c_int(0), C.c_int(0)
        n = lib.vsref_camera_parameters(h, buf.ctypes.data, 64, C.byref(w), C.byref(hh))
        return mt, w.value, hh.value, buf[:n].copy()
    finally:
        lib.vsref_destroy(h)


class CompiledRefTracker:
    """The reference's compiled ``frontend::FeatureTracker`` behind the oracle's tracker interface."""

    def __init__(self, cfg: RefConfig, camera: cam.Camera, per_instance_ids: bool = True, fisheye_mask_path: str = ""):
        self.lib = load()
        self.cfg = cfg
        self.m_camera = camera
        self.fisheye_mask_path = fisheye_mask_path
        self.find_f = None            # optional override: fn(pts1, pts2) -> (F or None, status or None)
        self.stage = {}
        self._n_id = 0
        self.per_instance_ids = per_instance_ids
        self.h = self.lib.vsref_create()
        self._set_camera(camera)
        self._pull()

    def _set_camera(self, c: cam.Camera):
        w, h = int(self.cfg.col), int(self.cfg.row)
        if c.model in (cam.PINHOLE, cam.KANNALA_BRANDT):
            # FeatureTracker::setIntrinsicParameter (ft:322-336): k2..k5 are k1,k2,p1,p2 for PINHOLE
            self.lib.vsref_set_intrinsic(self.h, c.model, w, h, c.fx, c.fy, c.cx, c.cy, *c.d)
        elif c.model == cam.MEI:
            self.lib.vsref_set_camera_mei(self.h, w, h, c.xi, *c.d, c.fx, c.fy, c.cx, c.cy)
        elif c.model == cam.SCARAMUZZA:
            poly = (C.c_double * 5)(*c.poly)
            self.lib.vsref_set_camera_ocam(self.h, w, h, c.C, c.D, c.E, c.cx, c.cy, poly)
        else:
            raise ValueError(c.model)

    def close(self):
        if self.h:
            self.lib.vsref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _enter(self):
        global _active, _cb_error
        c = self.cfg
        cc = _Config(float(c.row), float(c.col), float(c.focal_length), int(c.max_cnt), int(c.min_dist),
                     float(c.f_threshold), int(c.equalize), int(c.fisheye), int(c.lk_window_size),
                     int(c.lk_pyramid_levels), int(c.lk_iterations), float(c.lk_eps), float(c.f_threshold_edge_factor))
        self.lib.vsref_set_config(C.byref(cc), self.fisheye_mask_path.encode())
        if self.per_instance_ids:
            self.lib.vsref_set_n_id(self._n_id)
        _active, _cb_error = self, None

    def _leave(self):
        global _active, _cb_error
        _active = None
        if self.per_instance_ids:
            self._n_id = self.lib.vsref_get_n_id()
        if _cb_error is not None:
            e, _cb_error = _cb_error, None
            raise e

    def _pull(self):
        n = self.lib.vsref_count(self.h)
        ids = np.zeros(n, np.int32)
        cnt = np.zeros(n, np.int32)
        pts = np.zeros((n, 2), np.float32)
        un = np.zeros((n, 2), np.float32)
        vel = np.zeros((n, 2), np.float32)
        if n:
            self.lib.vsref_get(self.h, ids.ctypes.data, cnt.ctypes.data, pts.ctypes.data, un.ctypes.data, vel.ctypes.data)
        self.ids, self.track_cnt = ids.tolist(), cnt.tolist()
        self.cur_pts, self.cur_undistorted_pts, self.pts_velocity = pts, un, vel

    def detect_and_track(self, img: np.ndarray, t: float):
        assert img.dtype == np.uint8 and img.ndim == 2 and img.strides[1] == 1
        self.stage = {"ransac_ran": False}
        self._enter()
        try:
            self.lib.vsref_detect_and_track(self.h, img.ctypes.data, img.shape[0], img.shape[1], img.strides[0], float(t))
        finally:
            self._leave()
        self._pull()
        self.n_new = self.lib.vsref_n_new(self.h)

    def update_all_ids(self):
        self._enter()
        try:
            self.lib.vsref_update_ids(self.h)
        finally:
            self._leave()
        self._pull()

    @property
    def mask(self):
        m = np.zeros((int(self.cfg.row), int(self.cfg.col)), np.uint8)
        return m if self.lib.vsref_get_mask(self.h, m.ctypes.data) else None

    def lift_batch(self, pts) -> np.ndarray:
        """Camera::liftProjective of the compiled reference camera over (N,2) pixels -> (N,3) float64."""
        uv = np.ascontiguousarray(pts, np.float64).reshape(-1, 2)
        out = np.zeros((len(uv), 3))
        if len(uv):
            self._enter()
            try:
                self.lib.vsref_lift_projective(self.h, len(uv), uv.ctypes.data, out.ctypes.data)
            finally:
                self._leave()
        return out

    def in_border(self, pt) -> bool:
        self._enter()
        try:
            return bool(self.lib.vsref_in_border(float(pt[0]), float(pt[1])))
        finally:
            self._leave()

    def published(self):
        """What the callers forward: only track_cnt > 1 (measurement_processor.cpp:212-223)."""
        out = {}
        for j in range(len(self.ids)):
            if self.track_cnt[j] > 1:
                out[self.ids[j]] = (float(self.cur_undistorted_pts[j][0]), float(self.cur_undistorted_pts[j][1]),
                                    1.0, float(self.cur_pts[j][0]), float(self.cur_pts[j][1]),
                                    float(self.pts_velocity[j][0]), float(self.pts_velocity[j][1]))
        return out
