"""Frame ingest — the step before the path, host side of ``csrc/ingest.cu``.

``MeasurementProcessor::extractImageFeatures`` (``/root/reference/src/utility/measurement_processor.cpp:178-231``) reads
each frame with ``cv::imread(path, IMREAD_GRAYSCALE)``, tracks ``rowRange(0, camera.row)`` of it, runs the ``updateID``
loop and publishes the ``track_cnt > 1`` features; ``VIOEngine::processFrame`` (``src/vio_engine.cpp:180-239``) does the
same from a raw grey pointer.  ``IngestPipeline`` is that caller for a batch of streams: files (or raw frames) in,
published records out, with decode / upload / kernels overlapped across three batches.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi

__all__ = ["decode_gray", "read_gray", "IngestPipeline"]


def decode_gray(data: bytes, max_w: int = 16384, max_h: int = 16384) -> np.ndarray:
    """PNG / binary-PGM bytes -> (rows, W) uint8, exactly what ``cv::imread(..., IMREAD_GRAYSCALE)`` returns, cropped to
    the first ``max_h`` rows (the reference's ``rowRange(0, camera.row)``)."""
    lib = _abi.load()
    w, h = C.c_int32(0), C.c_int32(0)
    # probe the size first with a one-row buffer, then decode into a buffer of the real size
    probe = np.zeros((1, max_w), np.uint8)
    buf = (C.c_ubyte * len(data)).from_buffer_copy(data)
    _abi.check(lib.vs_image_decode_gray(buf, len(data), probe.ctypes.data, max_w, max_w, 1, C.byref(w), C.byref(h)))
    rows = min(h.value, max_h)
    out = np.zeros((rows, w.value), np.uint8)
    _abi.check(lib.vs_image_decode_gray(buf, len(data), out.ctypes.data, w.value, w.value, rows, C.byref(w), C.byref(h)))
    return out


def read_gray(path: str, max_w: int = 16384, max_h: int = 16384) -> np.ndarray:
    with open(path, "rb") as f:
        return decode_gray(f.read(), max_w, max_h)


class IngestPipeline:
    """files / raw frames -> pinned ring -> ``vs_tracker_submit_batch`` / ``vs_tracker_wait`` around one ``BatchTracker``.

    While the pipeline exists the tracker must not be driven directly.  At most three batches may be outstanding
    (``submit_*`` without a matching ``wait``)."""

    def __init__(self, tracker, decode_threads: int = 4):
        self._lib = _abi.load()
        self._bt = tracker
        self._h = C.c_void_p()
        _abi.check(self._lib.vs_ingest_create(tracker._h, int(decode_threads), C.byref(self._h)))
        self._res = (_abi.VsTrackResult * tracker.n_streams)()

    def close(self):
        if self._h:
            self._lib.vs_ingest_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def submit_files(self, paths, times):
        n = self._bt.n_streams
        assert len(paths) == n and len(times) == n
        arr = (C.c_char_p * n)(*[str(p).encode() for p in paths])
        ts = (C.c_double * n)(*[float(t) for t in times])
        _abi.check(self._lib.vs_ingest_submit_files(self._h, arr, ts))

    def submit_raw(self, frames, times):
        """frames: n_streams (H, W) uint8 arrays; copied into the pinned ring before this returns (the reference's raw
        pointer is only valid during the call, vio_engine.cpp:180)."""
        n = self._bt.n_streams
        fr = [np.ascontiguousarray(f, np.uint8) for f in frames]
        ptrs = (C.c_void_p * n)(*[f.ctypes.data for f in fr])
        ts = (C.c_double * n)(*[float(t) for t in times])
        _abi.check(self._lib.vs_ingest_submit_raw(self._h, ptrs, self._bt.tc.col, ts))

    def wait(self):
        """Results of the OLDEST outstanding batch (list of StreamResult, one per stream)."""
        _abi.check(self._lib.vs_ingest_wait(self._h, self._res))
        saved = self._bt._res
        self._bt._res = self._res
        try:
            return self._bt._collect()
        finally:
            self._bt._res = saved
