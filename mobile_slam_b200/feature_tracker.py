"""Host-side mirror of ``frontend::FeatureTracker`` (reference include/frontend/feature_tracker.h:25-76)
over the C ABI of ``libvs_tracker.so``.

``FeatureTracker``   one camera stream, the reference's method names and public vectors
                     (detectAndTrack / readImage alias / updateID / setIntrinsicParameter /
                     readIntrinsicParameter; cur_pts, cur_undistorted_pts, pts_velocity, ids, track_cnt).
``BatchTracker``     n independent streams tracked in lock-step by one handle (stream-parallel batching,
                     SURVEY.md §8e).

All computation happens in CUDA kernels behind the ABI; this file only marshals buffers.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _abi


@dataclass
class TrackerConfig:
    """utility::g_config.camera / .feature_tracker fields the tracker reads (include/utility/config.h:18-56)."""
    col: int = 752
    row: int = 480
    focal_length: float = 460.0
    max_cnt: int = 150
    min_dist: int = 30
    f_threshold: float = 1.0
    f_threshold_edge_factor: float = 0.0
    equalize: int = 1
    fisheye: int = 0
    lk_window_size: int = 21
    lk_pyramid_levels: int = 3
    lk_iterations: int = 30
    lk_eps: float = 0.01
    fb_check: int = 0
    fb_threshold: float = 0.5
    ransac_mode: int = _abi.VS_RANSAC_DEVICE
    capacity: int = 0
    device: int = 0
    first_feature_id: int = 0
    pipeline_lanes: int = 0
    fisheye_mask: np.ndarray | None = None


@dataclass
class CameraParams:
    model_type: int = _abi.VS_CAM_PINHOLE
    width: int = 752
    height: int = 480
    fx: float = 460.0
    fy: float = 460.0
    cx: float = 376.0
    cy: float = 240.0
    d: tuple = (0.0, 0.0, 0.0, 0.0)
    xi: float = 0.0
    poly: tuple = (0.0,) * 5
    affine: tuple = (1.0, 0.0, 0.0)


def config_from_synth(cfg: dict, **over):
    """(TrackerConfig, CameraParams) from a ``synth.CONFIGS`` entry."""
    tc = TrackerConfig(col=cfg["width"], row=cfg["height"], focal_length=cfg["focal_length"],
                       max_cnt=cfg["max_cnt"], min_dist=cfg["min_dist"], fisheye=cfg["fisheye"],
                       lk_pyramid_levels=cfg["max_level"])
    for k, v in over.items():
        setattr(tc, k, v)
    fx, fy, cx, cy = cfg["intr"]
    cam = CameraParams(_abi.CAMERA_MODELS[cfg["model"]], cfg["width"], cfg["height"], fx, fy, cx, cy,
                       tuple(cfg["dist"]), cfg.get("xi", 0.0), tuple(cfg.get("poly", (0.0,) * 5)),
                       tuple(cfg.get("affine", (1.0, 0.0, 0.0))))
    return tc, cam


def _fill_config(tc: TrackerConfig, cam: CameraParams, n_streams: int):
    lib = _abi.load()
    c = _abi.VsTrackerConfig()
    lib.vs_tracker_default_config(C.byref(c))
    c.device, c.n_streams = tc.device, n_streams
    c.width, c.height, c.focal_length = tc.col, tc.row, tc.focal_length
    c.max_cnt, c.min_dist = tc.max_cnt, tc.min_dist
    c.f_threshold, c.f_threshold_edge_factor = tc.f_threshold, tc.f_threshold_edge_factor
    c.equalize, c.fisheye = tc.equalize, tc.fisheye
    c.lk_window_size, c.lk_pyramid_levels = tc.lk_window_size, tc.lk_pyramid_levels
    c.lk_iterations, c.lk_eps = tc.lk_iterations, tc.lk_eps
    c.fb_check, c.fb_threshold = tc.fb_check, tc.fb_threshold
    c.ransac_mode, c.capacity, c.first_feature_id = tc.ransac_mode, tc.capacity, tc.first_feature_id
    c.pipeline_lanes = tc.pipeline_lanes
    keep = None
    if tc.fisheye_mask is not None:
        keep = np.ascontiguousarray(tc.fisheye_mask, np.uint8)
        if keep.shape != (tc.row, tc.col):
            raise ValueError("fisheye_mask must be (row, col) uint8")
        c.fisheye_mask_image = keep.ctypes.data
        c.fisheye_mask_stride = tc.col
    k = c.camera
    k.model_type, k.width, k.height = cam.model_type, cam.width, cam.height
    k.fx, k.fy, k.cx, k.cy = cam.fx, cam.fy, cam.cx, cam.cy
    for i in range(4):
        k.d[i] = cam.d[i]
    k.xi = cam.xi
    for i in range(5):
        k.poly[i] = cam.poly[i]
    k.aff_c, k.aff_d, k.aff_e = cam.affine
    return c, keep


@dataclass
class StreamResult:
    ids: np.ndarray
    track_cnt: np.ndarray
    cur_pts: np.ndarray
    cur_undistorted_pts: np.ndarray
    pts_velocity: np.ndarray
    n_new: int
    counters: dict = field(default_factory=dict)


class BatchTracker:
    """n_streams independent FeatureTrackers sharing every kernel launch."""

    def __init__(self, tc: TrackerConfig, cam: CameraParams, n_streams: int = 1):
        self._lib = _abi.load()
        self.tc, self.cam, self.n_streams = tc, cam, int(n_streams)
        cfg, keep = _fill_config(tc, cam, self.n_streams)
        eff = _abi.VsTrackerConfig()
        self._h = C.c_void_p()
        _abi.check(self._lib.vs_tracker_create(C.byref(cfg), C.byref(eff), C.byref(self._h)))
        self.capacity = eff.capacity
        self._res = (_abi.VsTrackResult * self.n_streams)()
        self._cb = None

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.vs_tracker_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- frame path -------------------------------------------------------------------------
    def _collect(self):
        out = []
        for s in range(self.n_streams):
            r = self._res[s]
            n = r.n

            def arr(p, k, dt):
                return np.ctypeslib.as_array(p, shape=(n * k,)).astype(dt, copy=True).reshape((n, k) if k > 1 else (n,)) \
                    if n else np.zeros((0, k) if k > 1 else (0,), dt)
            out.append(StreamResult(arr(r.ids, 1, np.int32), arr(r.track_cnt, 1, np.int32),
                                    arr(r.cur_pts, 2, np.float32), arr(r.undist, 2, np.float32),
                                    arr(r.velocity, 2, np.float32), r.n_new,
                                    dict(n_lk_in=r.n_lk_in, n_lk_ok=r.n_lk_ok, n_ransac_in=r.n_ransac_in,
                                         n_ransac_out=r.n_ransac_out, n_mask_kept=r.n_mask_kept,
                                         n_candidates=r.n_candidates, ransac_ran=r.ransac_ran,
                                         rms_disp=r.rms_disp)))
        return out

    def track(self, frames, times, collect: bool = True):
        """frames: sequence of n_streams (H, W) uint8 host arrays; times: n_streams floats."""
        if len(frames) != self.n_streams:
            raise ValueError("one frame per stream expected")
        # row-strided views (a cv::Mat with step > cols, vio_engine.cpp:180 wraps the caller's buffer without a copy)
        # go to the library as they are; anything else is made contiguous first
        steps = {f.strides[0] for f in frames
                 if isinstance(f, np.ndarray) and f.dtype == np.uint8 and f.ndim == 2 and f.strides[1] == 1}
        if len(steps) == 1 and all(isinstance(f, np.ndarray) and f.dtype == np.uint8 and f.ndim == 2
                                   and f.strides[1] == 1 for f in frames):
            fr, stride = list(frames), steps.pop()
        else:
            fr, stride = [np.ascontiguousarray(f, np.uint8) for f in frames], self.tc.col
        for f in fr:
            if f.shape != (self.tc.row, self.tc.col):
                raise ValueError(f"frame shape {f.shape} != {(self.tc.row, self.tc.col)}")
        ptrs = (C.c_void_p * self.n_streams)(*[f.ctypes.data for f in fr])
        ts = (C.c_double * self.n_streams)(*[float(t) for t in times])
        _abi.check(self._lib.vs_tracker_track_batch(self._h, ptrs, stride, ts, self._res))
        return self._collect() if collect else None

    def track_pinned(self, base_ptr: int, image_stride: int, times_ctypes, collect: bool = False):
        """Frames packed in one (pinned) host block: stream s at base_ptr + s*image_stride."""
        ptrs = (C.c_void_p * self.n_streams)(*[base_ptr + s * image_stride for s in range(self.n_streams)])
        _abi.check(self._lib.vs_tracker_track_batch(self._h, ptrs, self.tc.col, times_ctypes, self._res))
        return self._collect() if collect else None

    def submit_pinned(self, base_ptr: int, image_stride: int, times_ctypes):
        """Asynchronous track_pinned: returns at once; pair with wait() (at most two submits in flight)."""
        ptrs = (C.c_void_p * self.n_streams)(*[base_ptr + s * image_stride for s in range(self.n_streams)])
        _abi.check(self._lib.vs_tracker_submit_batch(self._h, ptrs, self.tc.col, times_ctypes))

    def submit(self, frames, times):
        fr = [np.ascontiguousarray(f, np.uint8) for f in frames]
        self._inflight = getattr(self, "_inflight", []) + [fr]          # keep the buffers alive until wait()
        ptrs = (C.c_void_p * self.n_streams)(*[f.ctypes.data for f in fr])
        ts = (C.c_double * self.n_streams)(*[float(t) for t in times])
        _abi.check(self._lib.vs_tracker_submit_batch(self._h, ptrs, self.tc.col, ts))

    def wait(self, collect: bool = True):
        _abi.check(self._lib.vs_tracker_wait(self._h, self._res))
        if getattr(self, "_inflight", None):
            self._inflight.pop(0)
        return self._collect() if collect else None

    def track_device(self, d_ptr: int, stride: int, image_stride: int, times, to_host: bool = True,
                     collect: bool = True):
        ts = times if isinstance(times, C.Array) else (C.c_double * self.n_streams)(*[float(t) for t in times])
        _abi.check(self._lib.vs_tracker_track_batch_device(self._h, C.c_void_p(d_ptr), stride, image_stride, ts,
                                                            1 if to_host else 0, self._res))
        return self._collect() if (to_host and collect) else None

    def fetch(self):
        _abi.check(self._lib.vs_tracker_fetch_results(self._h, self._res))
        return self._collect()

    def set_params(self, max_cnt=-1, min_dist=-1, f_threshold=-1.0, lk_window=-1, lk_pyramid=-1,
                   lk_iterations=-1, lk_eps=-1.0, f_edge_factor=-1.0):
        _abi.check(self._lib.vs_tracker_set_params(self._h, max_cnt, min_dist, f_threshold, lk_window, lk_pyramid,
                                                   lk_iterations, lk_eps, f_edge_factor))

    def set_frontend_flags(self, equalize=-1, fisheye=-1, focal_length=-1.0, fisheye_mask=None):
        """g_config.feature_tracker.equalize / .fisheye (+ mask image) / camera.focal_length, which the reference re-reads
        on every call (feature_tracker.cpp:39-46,102,214-222); tracking state is kept."""
        m = None if fisheye_mask is None else np.ascontiguousarray(fisheye_mask, np.uint8)
        _abi.check(self._lib.vs_tracker_set_frontend_flags(self._h, int(equalize), int(fisheye), float(focal_length),
                                                           None if m is None else m.ctypes.data,
                                                           0 if m is None else m.strides[0]))

    def set_reject_callback(self, fn):
        """fn(stream, un_cur (n,2) f32, un_next (n,2) f32, next_px (n,2) f32) -> status (n,) {0,1}."""
        if fn is None:
            self._cb = None
            _abi.check(self._lib.vs_tracker_set_reject_callback(self._h, _abi.REJECT_FN(), None))
            return

        def tramp(_user, stream, n, a, b, px, status):
            try:
                A = np.ctypeslib.as_array(a, shape=(n, 2)).copy()
                Bm = np.ctypeslib.as_array(b, shape=(n, 2)).copy()
                P = np.ctypeslib.as_array(px, shape=(n, 2)).copy()
                st = np.asarray(fn(stream, A, Bm, P), np.uint8).reshape(-1)
                np.ctypeslib.as_array(status, shape=(n,))[:] = st[:n]
                return 0
            except Exception:  # noqa: BLE001 - surfaced as a C error code
                import traceback
                traceback.print_exc()
                return 1
        self._cb = _abi.REJECT_FN(tramp)
        _abi.check(self._lib.vs_tracker_set_reject_callback(self._h, self._cb, None))

    def reset(self, stream: int = -1):
        _abi.check(self._lib.vs_tracker_reset(self._h, stream))

    def debug_guard_check(self) -> int:
        """Overwritten guard bytes around the handle's device buffers (needs VS_DEBUG_GUARD=1 at creation; -1 without)."""
        return int(self._lib.vs_tracker_debug_guard_check(self._h))

    def debug_read_pyramid(self, stream: int = 0):
        """[(level incl. its stored border, pad_x, pad_y)] of the newest frame of ``stream``, as the LK kernel sees it."""
        layout = (C.c_int64 * 27)()
        nl = C.c_int32(0)
        _abi.check(self._lib.vs_tracker_debug_read_pyramid(self._h, stream, None, 0, layout, C.byref(nl)))
        buf = np.zeros(layout[24], np.uint8)
        _abi.check(self._lib.vs_tracker_debug_read_pyramid(self._h, stream, buf.ctypes.data, buf.size, layout, C.byref(nl)))
        px, py = int(layout[25]), int(layout[26])
        out = []
        for l in range(nl.value):
            off, pitch, w, h = (int(layout[4 * l + k]) for k in range(4))
            start = off - py * pitch - px
            block = buf[start:start + (h + 2 * py) * pitch].reshape(h + 2 * py, pitch)
            out.append((block[:, :w + 2 * px].copy(), px, py))
        return out

    @property
    def launch_count(self) -> int:
        return int(self._lib.vs_tracker_launch_count(self._h))

    @property
    def last_device_ms(self) -> float:
        return float(self._lib.vs_tracker_last_device_ms(self._h))

    # ---- stage entry points (teacher-forced parity tests; they reset tracking state) ----------
    def stage_clahe(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        out = np.empty_like(img)
        _abi.check(self._lib.vs_stage_clahe(self._h, img.ctypes.data, out.ctypes.data))
        return out

    def stage_pyramid(self, level0):
        level0 = np.ascontiguousarray(level0, np.uint8)
        buf = np.empty(level0.size * 2, np.uint8)
        sizes = np.zeros(16, np.int32)
        nl = C.c_int32(0)
        _abi.check(self._lib.vs_stage_pyramid(self._h, level0.ctypes.data, buf.ctypes.data, sizes.ctypes.data, C.byref(nl)))
        out, off = [], 0
        for l in range(nl.value):
            w, h = int(sizes[2 * l]), int(sizes[2 * l + 1])
            out.append(buf[off:off + w * h].reshape(h, w).copy())
            off += w * h
        return out

    def stage_lk(self, prev0, next0, pts):
        prev0 = np.ascontiguousarray(prev0, np.uint8)
        next0 = np.ascontiguousarray(next0, np.uint8)
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
        n = len(pts)
        nxt = np.zeros((n, 2), np.float32)
        st = np.zeros(n, np.uint8)
        _abi.check(self._lib.vs_stage_lk(self._h, prev0.ctypes.data, next0.ctypes.data, pts.ctypes.data, n,
                                         nxt.ctypes.data, st.ctypes.data))
        return nxt, st

    def stage_min_eig(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        eig = np.zeros(img.shape, np.float32)
        _abi.check(self._lib.vs_stage_min_eig(self._h, img.ctypes.data, eig.ctypes.data))
        return eig

    def stage_good_features(self, img, kept_pts, max_corners, return_mask=False):
        img = np.ascontiguousarray(img, np.uint8)
        kept = np.ascontiguousarray(kept_pts, np.float32).reshape(-1, 2)
        out = np.zeros((max(max_corners, 1), 2), np.float32)
        n = C.c_int32(0)
        mask = np.zeros(img.shape, np.uint8) if return_mask else None
        _abi.check(self._lib.vs_stage_good_features(self._h, img.ctypes.data, kept.ctypes.data if len(kept) else None,
                                                    len(kept), max_corners, out.ctypes.data, C.byref(n),
                                                    mask.ctypes.data if return_mask else None))
        return (out[:n.value].copy(), mask) if return_mask else out[:n.value].copy()

    def stage_undistort(self, pts):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
        out = np.zeros_like(pts)
        _abi.check(self._lib.vs_stage_undistort(self._h, pts.ctypes.data, len(pts), out.ctypes.data))
        return out

    def stage_set_mask(self, pts, track_cnt):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
        tc = np.ascontiguousarray(track_cnt, np.int32)
        keep = np.zeros(max(len(pts), 1), np.int32)
        n = C.c_int32(0)
        _abi.check(self._lib.vs_stage_set_mask(self._h, pts.ctypes.data, tc.ctypes.data, len(pts), keep.ctypes.data,
                                               C.byref(n)))
        return keep[:n.value].copy()

    def stage_ransac(self, a, b):
        a = np.ascontiguousarray(a, np.float32).reshape(-1, 2)
        b = np.ascontiguousarray(b, np.float32).reshape(-1, 2)
        st = np.zeros(len(a), np.uint8)
        _abi.check(self._lib.vs_stage_ransac(self._h, a.ctypes.data, b.ctypes.data, len(a), st.ctypes.data))
        return st

    # ---- measurement hooks -------------------------------------------------------------------
    def stage_ransac_mark(self, F, a, b, next_px, edge_factor=0.0, base_status=None):
        """flags the device assigns for a GIVEN fundamental matrix (teacher forcing); base_status = a teacher's inlier
        mask -> only the edge-recovery rule of feature_tracker.cpp:236-285 is applied on top of it"""
        F = np.ascontiguousarray(F, np.float64).reshape(9)
        a = np.ascontiguousarray(a, np.float32).reshape(-1, 2)
        b = np.ascontiguousarray(b, np.float32).reshape(-1, 2)
        px = np.ascontiguousarray(next_px, np.float32).reshape(-1, 2)
        st = np.zeros(len(a), np.uint8)
        base = None if base_status is None else np.ascontiguousarray(base_status, np.uint8)
        _abi.check(self._lib.vs_stage_ransac_mark(self._h, F.ctypes.data, a.ctypes.data, b.ctypes.data, px.ctypes.data, len(a),
                                                  float(edge_factor), None if base is None else base.ctypes.data, st.ctypes.data))
        return st

    @property
    def cuda_stream(self) -> int:
        """cudaStream_t of the handle (wrap with torch.cuda.ExternalStream to record events on it)."""
        return int(self._lib.vs_tracker_cuda_stream(self._h) or 0)

    def set_profiling(self, on: bool):
        _abi.check(self._lib.vs_tracker_set_profiling(self._h, 1 if on else 0))

    def profile(self) -> dict:
        """{kernel name: (total_ms, launches)} accumulated since set_profiling(True)."""
        ms = (C.c_float * 32)()
        cnt = (C.c_int32 * 32)()
        n = C.c_int32(0)
        _abi.check(self._lib.vs_tracker_get_profile(self._h, 32, ms, cnt, C.byref(n)))
        return {self._lib.vs_profile_kernel_name(i).decode(): (float(ms[i]), int(cnt[i])) for i in range(n.value)}


CLAHE_NO_ALLOC_MAX_IMAGES = 64      # VS_CLAHE_NO_ALLOC_MAX_IMAGES (include/vs_tracker.h)


def clahe_no_alloc(d_src: int, src_pitch: int, src_image_stride: int, d_dst: int, dst_pitch: int, dst_image_stride: int,
                   width: int, height: int, n_images: int = 1, cuda_stream: int = 0) -> None:
    """``vs_clahe_no_alloc``: CLAHE of ``n_images`` device images (raw device addresses) enqueued on ``cuda_stream`` without
    any allocation — no handle needed; the scratch is a static device array shared by every call on the device (calls
    that overlap on the GPU race; at most CLAHE_NO_ALLOC_MAX_IMAGES images per call).  Not synchronised."""
    _abi.check(_abi.load().vs_clahe_no_alloc(d_src, src_pitch, src_image_stride, d_dst, dst_pitch, dst_image_stride,
                                             width, height, n_images, cuda_stream or None))


class FeatureTracker:
    """Drop-in mirror of ``frontend::FeatureTracker`` for one camera stream.

    Reads its parameters from a ``TrackerConfig`` (the stand-in for the global ``utility::g_config``)
    when the first frame arrives, like the reference reads g_config on every call.
    """

    n_id = 0   # static int FeatureTracker::n_id (feature_tracker.cpp:10); ids are per stream here

    def __init__(self, config: TrackerConfig | None = None):
        self.config = config or TrackerConfig()
        self._cam: CameraParams | None = None
        self._bt: BatchTracker | None = None
        self.cur_pts = np.zeros((0, 2), np.float32)
        self.cur_undistorted_pts = np.zeros((0, 2), np.float32)
        self.pts_velocity = np.zeros((0, 2), np.float32)
        self.ids: list[int] = []
        self.track_cnt: list[int] = []
        self._real_ids: list[int] = []
        self.cur_time = 0.0
        self.prev_time = 0.0
        self.counters: dict = {}

    # feature_tracker.cpp:322-336 (note the enum trap: anything but PINHOLE(=2) is equidistant)
    def setIntrinsicParameter(self, model_type, width, height, fx, fy, cx, cy, k2, k3, k4, k5):
        mt = _abi.VS_CAM_PINHOLE if model_type == _abi.VS_CAM_PINHOLE else _abi.VS_CAM_KANNALA_BRANDT
        self._cam = CameraParams(mt, width, height, fx, fy, cx, cy, (k2, k3, k4, k5))

    # feature_tracker.cpp:318-320 -> CameraFactory::generateCameraFromYamlFile (CameraFactory.cc:87-150)
    def readIntrinsicParameter(self, calib_file: str):
        from .calib import camera_from_yaml
        self._cam = camera_from_yaml(calib_file)

    def setCamera(self, cam: CameraParams):
        self._cam = cam

    _RUNTIME = ("max_cnt", "min_dist", "f_threshold", "lk_window_size", "lk_pyramid_levels", "lk_iterations", "lk_eps",
                "f_threshold_edge_factor")
    _FLAGS = ("equalize", "fisheye", "focal_length")

    def _ensure(self):
        if self._bt is None:
            if self._cam is None:
                raise RuntimeError("camera intrinsics not set (readIntrinsicParameter/setIntrinsicParameter)")
            self._bt = BatchTracker(self.config, self._cam, 1)
            self._applied = {k: getattr(self.config, k) for k in self._RUNTIME + self._FLAGS}

    def _apply_runtime_config(self):
        """The reference reads ``g_config`` inside every call (feature_tracker.cpp:39-46,102,118-119,132-134,214-222):
        changes to ``self.config`` between frames take effect on the next frame, tracking state kept."""
        c, a = self.config, self._applied
        if any(getattr(c, k) != a[k] for k in self._RUNTIME):
            self._bt.set_params(c.max_cnt, c.min_dist, c.f_threshold, c.lk_window_size, c.lk_pyramid_levels,
                                c.lk_iterations, c.lk_eps, c.f_threshold_edge_factor)
        if any(getattr(c, k) != a[k] for k in self._FLAGS):
            mask = c.fisheye_mask if (c.fisheye and not a["fisheye"]) else None
            self._bt.set_frontend_flags(1 if c.equalize else 0, 1 if c.fisheye else 0, c.focal_length, mask)
        self._applied = {k: getattr(c, k) for k in self._RUNTIME + self._FLAGS}

    def detectAndTrack(self, img, cur_time: float):
        self._ensure()
        img = np.asarray(img)
        if img.shape != (self.config.row, self.config.col):
            # OpenCV throws in calcOpticalFlowPyrLK when the frame size changes; the handle's buffers are sized once
            raise ValueError(f"image is {img.shape[1]}x{img.shape[0]}, camera.col x camera.row is {self.config.col}x{self.config.row}")
        self._apply_runtime_config()
        self.cur_time = float(cur_time)
        r = self._bt.track([img], [cur_time])[0]
        self.cur_pts = r.cur_pts
        self.cur_undistorted_pts = r.cur_undistorted_pts
        self.pts_velocity = r.pts_velocity
        self.track_cnt = r.track_cnt.tolist()
        self._real_ids = r.ids.tolist()
        n_old = len(self._real_ids) - r.n_new
        # the reference leaves new points at id -1 until the caller's updateID loop runs
        self.ids = self._real_ids[:n_old] + [-1] * r.n_new
        self.counters = r.counters
        self.prev_time = self.cur_time

    readImage = detectAndTrack   # upstream VINS-Mono's name for the same entry point

    # feature_tracker.cpp:309-316
    def updateID(self, i: int) -> bool:
        if i < len(self.ids):
            if self.ids[i] == -1:
                self.ids[i] = self._real_ids[i]
                FeatureTracker.n_id = max(FeatureTracker.n_id, self._real_ids[i] + 1)
            return True
        return False

    def image_data(self) -> dict:
        """common::ImageData the callers build: id -> [x, y, 1, u, v, vx, vy] for track_cnt > 1
        (measurement_processor.cpp:212-223, vio_engine.cpp:219-239)."""
        out = {}
        for j, pid in enumerate(self.ids):
            if self.track_cnt[j] > 1:
                out[pid] = (float(self.cur_undistorted_pts[j][0]), float(self.cur_undistorted_pts[j][1]), 1.0,
                            float(self.cur_pts[j][0]), float(self.cur_pts[j][1]),
                            float(self.pts_velocity[j][0]), float(self.pts_velocity[j][1]))
        return out
