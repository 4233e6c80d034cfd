"""Python mirror of the reference's ``frontend::FeatureManager`` over the device feature bank (``vs_bank_*`` of
``include/vs_tracker.h``; kernels in ``csrc/feature_bank.cu``).

``FeatureBank`` holds the banks of ``n_streams`` independent camera streams on one GPU; ``StreamFeatureManager`` is one
stream's view with the reference's method names (``/root/reference/include/frontend/feature_manager.h:70-92``), so a test
or an estimator written against ``FeatureManager`` reads the same.  There is no CPU path: the library must load and a
GPU must be present.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi

WINDOW_SIZE = 10  # VS_BANK_WINDOW_SIZE
MAX_OBS = WINDOW_SIZE + 1
OP_NONE, OP_REMOVE_BACK, OP_REMOVE_FRONT, OP_REMOVE_BACK_SHIFT_DEPTH, OP_REMOVE_FAILURES = range(5)

_pi32 = C.POINTER(C.c_int32)
_pf64 = C.POINTER(C.c_double)


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(_pi32)


def _f(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(_pf64)


class FeatureBank:
    def __init__(self, n_streams: int, capacity: int = 1024, device: int = 0, min_parallax: float = 10.0,
                 focal_length: float = 460.0, init_depth: float = 5.0):
        self._lib = _abi.load()
        self.B, self.capacity = n_streams, capacity
        self.min_parallax, self.focal_length, self.init_depth = min_parallax, focal_length, init_depth
        self._h = C.c_void_p()
        _abi.check(self._lib.vs_bank_create(device, n_streams, capacity, C.byref(self._h)))
        self.last_track_num = np.zeros(n_streams, np.int32)
        self.parallax_sum = np.zeros(n_streams)
        self.parallax_num = np.zeros(n_streams, np.int32)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.vs_bank_destroy(self._h)
            self._h = None

    __del__ = close

    def guard_check(self) -> int:
        """Guard bytes overwritten so far (needs VS_DEBUG_GUARD=1 at creation; -1 otherwise)."""
        return int(self._lib.vs_bank_debug_guard_check(self._h))

    def clear_state(self, stream: int = -1):
        _abi.check(self._lib.vs_bank_clear(self._h, stream))

    def add_frames(self, frame_count, images) -> np.ndarray:
        """addFeatureAndCheckParallax for every stream: ``images[s]`` = {feature_id: 7 doubles} (or None = no records).
        Returns the key-frame decisions [n_streams] (bool)."""
        B = self.B
        n = np.array([len(im) if im else 0 for im in images], dtype=np.int32)
        max_n = int(n.max()) if B else 0
        ids = np.zeros((B, max(max_n, 1)), np.int32)
        rec = np.zeros((B, max(max_n, 1), 7))
        for s, im in enumerate(images):
            if im:
                ids[s, :n[s]] = list(im.keys())
                rec[s, :n[s]] = np.array([im[k] for k in im], dtype=np.float64).reshape(-1, 7)
        fc, fcp = _i(np.broadcast_to(np.asarray(frame_count, dtype=np.int32), (B,)))
        n, np_ = _i(n)
        ids, idp = _i(ids)
        rec, recp = _f(rec)
        key = np.zeros(B, np.int32)
        _abi.check(self._lib.vs_bank_add_frame(self._h, fcp, np_, idp, recp, max_n, self.min_parallax, self.focal_length,
                                               key.ctypes.data_as(_pi32), self.last_track_num.ctypes.data_as(_pi32),
                                               self.parallax_sum.ctypes.data_as(_pf64), self.parallax_num.ctypes.data_as(_pi32)))
        return key.astype(bool)

    def add_frames_from_tracker(self, tracker, frame_count) -> np.ndarray:
        """The same, fed from the device-resident results of ``tracker``'s (a ``BatchTracker``) last frame."""
        fc, fcp = _i(np.broadcast_to(np.asarray(frame_count, dtype=np.int32), (self.B,)))
        key = np.zeros(self.B, np.int32)
        _abi.check(self._lib.vs_bank_add_frame_from_tracker(self._h, tracker._h, fcp, self.min_parallax, self.focal_length,
                                                            key.ctypes.data_as(_pi32), self.last_track_num.ctypes.data_as(_pi32)))
        return key.astype(bool)

    def slide(self, op, frame_count=None, poses=None):
        """op[s] in OP_*; ``poses[s]`` = (marg_R, marg_P, new_R, new_P) for OP_REMOVE_BACK_SHIFT_DEPTH."""
        op, opp = _i(np.broadcast_to(np.asarray(op, dtype=np.int32), (self.B,)))
        fcp = None
        if frame_count is not None:
            fc, fcp = _i(np.broadcast_to(np.asarray(frame_count, dtype=np.int32), (self.B,)))
        rpp = None
        if poses is not None:
            rp = np.zeros((self.B, 24))
            for s, q in enumerate(poses):
                if q is not None:
                    rp[s] = np.concatenate([np.asarray(q[0], dtype=np.float64).reshape(9), np.asarray(q[1], dtype=np.float64).reshape(3),
                                            np.asarray(q[2], dtype=np.float64).reshape(9), np.asarray(q[3], dtype=np.float64).reshape(3)])
            rp, rpp = _f(rp)
        _abi.check(self._lib.vs_bank_slide(self._h, opp, fcp, rpp, self.init_depth))

    def triangulate(self, poses, t_ic, r_ic):
        """``poses[s]`` = WINDOW_SIZE + 1 pairs (R, P), or None to leave stream s alone."""
        flat = np.zeros((self.B, MAX_OBS, 12))
        active, ap = _i([0 if w is None else 1 for w in poses])
        for s, w in enumerate(poses):
            for i, (R, P) in enumerate(w or []):
                flat[s, i, :9] = np.asarray(R, dtype=np.float64).reshape(9)
                flat[s, i, 9:] = np.asarray(P, dtype=np.float64)
        flat, fp = _f(flat)
        t, tp = _f(t_ic)
        r, rp = _f(np.asarray(r_ic, dtype=np.float64).reshape(9))
        _abi.check(self._lib.vs_bank_triangulate(self._h, fp, tp, rp, self.init_depth, ap))

    def get_depth_vector(self, stream: int) -> np.ndarray:
        out = np.zeros(self.capacity)
        n = C.c_int32(0)
        _abi.check(self._lib.vs_bank_get_depth_vector(self._h, stream, self.capacity, out.ctypes.data_as(_pf64), C.byref(n)))
        return out[:n.value].copy()

    def get_feature_count(self, stream: int) -> int:
        n = C.c_int32(0)
        _abi.check(self._lib.vs_bank_get_depth_vector(self._h, stream, 0, None, C.byref(n)))
        return n.value

    def set_depth(self, stream: int, x, clear: bool = False):
        x, xp = _f(x)
        _abi.check(self._lib.vs_bank_set_depth(self._h, stream, xp, len(x), 1 if clear else 0))

    def get_corresponding(self, stream: int, frame_l: int, frame_r: int) -> list:
        out = np.zeros((self.capacity, 6))
        n = C.c_int32(0)
        _abi.check(self._lib.vs_bank_get_corresponding(self._h, stream, frame_l, frame_r, self.capacity,
                                                       out.ctypes.data_as(_pf64), C.byref(n)))
        return [(list(out[i, :3]), list(out[i, 3:])) for i in range(n.value)]

    def export(self, stream: int):
        """(ids, start_frame, n_obs, depth, solve_flag, obs[n][MAX_OBS][7]) of one stream, in list order."""
        cap = self.capacity
        ids = np.zeros(cap, np.int32); start = np.zeros(cap, np.int32); n_obs = np.zeros(cap, np.int32)
        depth = np.zeros(cap); flag = np.zeros(cap, np.int32); obs = np.zeros((cap, MAX_OBS, 7))
        n = C.c_int32(0)
        _abi.check(self._lib.vs_bank_export(self._h, stream, cap, ids.ctypes.data_as(_pi32), start.ctypes.data_as(_pi32),
                                            n_obs.ctypes.data_as(_pi32), depth.ctypes.data_as(_pf64),
                                            flag.ctypes.data_as(_pi32), obs.ctypes.data_as(_pf64), C.byref(n)))
        k = n.value
        return ids[:k].copy(), start[:k].copy(), n_obs[:k].copy(), depth[:k].copy(), flag[:k].copy(), obs[:k].copy()


class StreamFeatureManager:
    """One stream of a ``FeatureBank`` behind the reference's ``FeatureManager`` method names (the other streams of the
    bank are left untouched by its calls)."""

    def __init__(self, bank: FeatureBank, stream: int = 0):
        self.bank, self.s = bank, stream
        self.last_track_num = 0

    def _per_stream(self, value, fill):
        v = [fill] * self.bank.B
        v[self.s] = value
        return v

    def clear_state(self):
        self.bank.clear_state(self.s)

    def add_feature_and_check_parallax(self, frame_count: int, image: dict) -> bool:
        key = self.bank.add_frames(frame_count, self._per_stream(image, None))
        self.last_track_num = int(self.bank.last_track_num[self.s])
        return bool(key[self.s])

    def get_feature_count(self) -> int:
        return self.bank.get_feature_count(self.s)

    def get_depth_vector(self) -> list:
        return list(self.bank.get_depth_vector(self.s))

    def set_depth(self, x):
        self.bank.set_depth(self.s, x, clear=False)

    def clear_depth(self, x):
        self.bank.set_depth(self.s, x, clear=True)

    def remove_failures(self):
        self.bank.slide(self._per_stream(OP_REMOVE_FAILURES, OP_NONE))

    def remove_back(self):
        self.bank.slide(self._per_stream(OP_REMOVE_BACK, OP_NONE))

    def remove_front(self, frame_count: int):
        self.bank.slide(self._per_stream(OP_REMOVE_FRONT, OP_NONE), frame_count)

    def remove_back_shift_depth(self, marg_R, marg_P, new_R, new_P):
        self.bank.slide(self._per_stream(OP_REMOVE_BACK_SHIFT_DEPTH, OP_NONE), None,
                        self._per_stream((marg_R, marg_P, new_R, new_P), None))

    def triangulate_across_all_views(self, poses, t_ic, r_ic):
        self.bank.triangulate(self._per_stream(poses, None), t_ic, r_ic)

    def get_corresponding(self, frame_l: int, frame_r: int) -> list:
        return self.bank.get_corresponding(self.s, frame_l, frame_r)

    def export(self):
        return self.bank.export(self.s)
