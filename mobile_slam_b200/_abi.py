"""ctypes binding of ``include/vs_tracker.h`` (the C ABI of libvs_tracker.so).

The shared library is built in-tree by ``mobile_slam_b200/csrc/build.sh`` (see ``__graft_entry__.build``).
There is no CPU fallback: if the library is missing, loading fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VS_TRACKER_LIB") or os.path.join(_HERE, "libvs_tracker.so")   # override: A/B builds of the same ABI

VS_OK = 0
VS_RANSAC_OFF, VS_RANSAC_DEVICE, VS_RANSAC_CALLBACK = 0, 1, 2
VS_CAM_KANNALA_BRANDT, VS_CAM_MEI, VS_CAM_PINHOLE, VS_CAM_SCARAMUZZA = 0, 1, 2, 3
CAMERA_MODELS = {"KANNALA_BRANDT": 0, "MEI": 1, "PINHOLE": 2, "SCARAMUZZA": 3}


class VsCamera(C.Structure):
    _fields_ = [("model_type", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
                ("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
                ("d", C.c_double * 4), ("xi", C.c_double), ("poly", C.c_double * 5),
                ("aff_c", C.c_double), ("aff_d", C.c_double), ("aff_e", C.c_double)]


class VsTrackerConfig(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("device", C.c_int32), ("n_streams", C.c_int32),
                ("width", C.c_int32), ("height", C.c_int32), ("focal_length", C.c_double),
                ("max_cnt", C.c_int32), ("min_dist", C.c_int32), ("f_threshold", C.c_double),
                ("f_threshold_edge_factor", C.c_double), ("equalize", C.c_int32), ("fisheye", C.c_int32),
                ("lk_window_size", C.c_int32), ("lk_pyramid_levels", C.c_int32),
                ("lk_iterations", C.c_int32), ("lk_eps", C.c_double), ("fb_check", C.c_int32),
                ("fb_threshold", C.c_double), ("ransac_mode", C.c_int32), ("capacity", C.c_int32),
                ("fisheye_mask_image", C.c_void_p), ("fisheye_mask_stride", C.c_int32),
                ("first_feature_id", C.c_int32), ("pipeline_lanes", C.c_int32), ("camera", VsCamera)]


class VsTrackResult(C.Structure):
    _fields_ = [("n", C.c_int32), ("n_new", C.c_int32),
                ("ids", C.POINTER(C.c_int32)), ("track_cnt", C.POINTER(C.c_int32)),
                ("cur_pts", C.POINTER(C.c_float)), ("undist", C.POINTER(C.c_float)),
                ("velocity", C.POINTER(C.c_float)),
                ("n_lk_in", C.c_int32), ("n_lk_ok", C.c_int32), ("n_ransac_in", C.c_int32),
                ("n_ransac_out", C.c_int32), ("n_mask_kept", C.c_int32), ("n_candidates", C.c_int32),
                ("ransac_ran", C.c_int32), ("rms_disp", C.c_float)]


REJECT_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_float),
                        C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_uint8))

# every symbol include/vs_tracker.h declares
EXPORTS = [
    "vs_last_error", "vs_abi_version", "vs_tracker_create", "vs_tracker_destroy",
    "vs_tracker_default_config", "vs_tracker_track", "vs_tracker_track_batch",
    "vs_tracker_track_batch_device", "vs_tracker_submit_batch", "vs_tracker_wait", "vs_tracker_fetch_results", "vs_tracker_set_params",
    "vs_tracker_set_reject_callback", "vs_tracker_reset", "vs_tracker_launch_count", "vs_tracker_debug_guard_check", "vs_clahe_no_alloc", "vs_tracker_debug_read_pyramid",
    "vs_tracker_last_device_ms", "vs_stage_clahe", "vs_stage_pyramid", "vs_stage_lk",
    "vs_stage_min_eig", "vs_stage_good_features", "vs_stage_undistort", "vs_stage_set_mask",
    "vs_stage_ransac", "vs_tracker_cuda_stream", "vs_tracker_set_profiling", "vs_tracker_get_profile",
    "vs_profile_kernel_name", "vs_pack_image_features", "vs_tracker_set_frontend_flags",
    "vs_tracker_get_config", "vs_image_decode_gray", "vs_image_read_gray", "vs_ingest_create", "vs_ingest_submit_files",
    "vs_ingest_submit_raw", "vs_ingest_wait", "vs_ingest_destroy", "vs_stage_ransac_mark",
    "vs_bank_create", "vs_bank_destroy", "vs_bank_clear", "vs_bank_add_frame", "vs_bank_add_frame_from_tracker",
    "vs_bank_slide", "vs_bank_triangulate", "vs_bank_get_depth_vector", "vs_bank_set_depth",
    "vs_bank_get_corresponding", "vs_bank_export", "vs_bank_debug_guard_check",
]

_lib = None


class VsError(RuntimeError):
    pass


def load():
    """Load libvs_tracker.so (raises if it has not been built — no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise VsError(f"{LIB_PATH} not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(mobile_slam_b200/csrc/build.sh). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    lib.vs_last_error.restype = C.c_char_p
    lib.vs_abi_version.restype = C.c_int
    lib.vs_tracker_create.argtypes = [C.POINTER(VsTrackerConfig), C.POINTER(VsTrackerConfig), C.POINTER(vp)]
    lib.vs_tracker_destroy.argtypes = [vp]
    lib.vs_tracker_destroy.restype = None
    lib.vs_tracker_default_config.argtypes = [C.POINTER(VsTrackerConfig)]
    lib.vs_tracker_default_config.restype = None
    lib.vs_tracker_track.argtypes = [vp, vp, i32, f64, C.POINTER(VsTrackResult)]
    lib.vs_tracker_track_batch.argtypes = [vp, C.POINTER(vp), i32, C.POINTER(f64), C.POINTER(VsTrackResult)]
    lib.vs_tracker_track_batch_device.argtypes = [vp, vp, i32, i64, C.POINTER(f64), i32, C.POINTER(VsTrackResult)]
    lib.vs_tracker_fetch_results.argtypes = [vp, C.POINTER(VsTrackResult)]
    lib.vs_tracker_submit_batch.argtypes = [vp, C.POINTER(vp), i32, C.POINTER(f64)]
    lib.vs_tracker_wait.argtypes = [vp, C.POINTER(VsTrackResult)]
    lib.vs_tracker_set_params.argtypes = [vp, i32, i32, f64, i32, i32, i32, f64, f64]
    lib.vs_tracker_set_frontend_flags.argtypes = [vp, i32, i32, f64, vp, i32]
    lib.vs_tracker_get_config.argtypes = [vp, C.POINTER(VsTrackerConfig)]
    lib.vs_image_decode_gray.argtypes = [vp, i64, vp, i32, i32, i32, C.POINTER(i32), C.POINTER(i32)]
    lib.vs_image_read_gray.argtypes = [C.c_char_p, vp, i32, i32, i32, C.POINTER(i32), C.POINTER(i32)]
    lib.vs_ingest_create.argtypes = [vp, i32, C.POINTER(vp)]
    lib.vs_ingest_submit_files.argtypes = [vp, C.POINTER(C.c_char_p), C.POINTER(f64)]
    lib.vs_ingest_submit_raw.argtypes = [vp, C.POINTER(vp), i32, C.POINTER(f64)]
    lib.vs_ingest_wait.argtypes = [vp, C.POINTER(VsTrackResult)]
    lib.vs_ingest_destroy.argtypes = [vp]
    lib.vs_ingest_destroy.restype = None
    lib.vs_tracker_set_reject_callback.argtypes = [vp, REJECT_FN, vp]
    lib.vs_tracker_reset.argtypes = [vp, i32]
    lib.vs_tracker_debug_guard_check.argtypes = [vp]
    lib.vs_tracker_debug_guard_check.restype = i64
    lib.vs_tracker_debug_read_pyramid.argtypes = [vp, i32, vp, i64, C.POINTER(i64), C.POINTER(i32)]
    lib.vs_clahe_no_alloc.argtypes = [vp, i32, i64, vp, i32, i64, i32, i32, i32, vp]
    lib.vs_tracker_launch_count.argtypes = [vp]
    lib.vs_tracker_launch_count.restype = i64
    lib.vs_tracker_last_device_ms.argtypes = [vp]
    lib.vs_tracker_last_device_ms.restype = C.c_float
    lib.vs_stage_clahe.argtypes = [vp, vp, vp]
    lib.vs_stage_pyramid.argtypes = [vp, vp, vp, vp, C.POINTER(i32)]
    lib.vs_stage_lk.argtypes = [vp, vp, vp, vp, i32, vp, vp]
    lib.vs_stage_min_eig.argtypes = [vp, vp, vp]
    lib.vs_stage_good_features.argtypes = [vp, vp, vp, i32, i32, vp, C.POINTER(i32), vp]
    lib.vs_stage_undistort.argtypes = [vp, vp, i32, vp]
    lib.vs_stage_set_mask.argtypes = [vp, vp, vp, i32, vp, C.POINTER(i32)]
    lib.vs_stage_ransac.argtypes = [vp, vp, vp, i32, vp]
    lib.vs_stage_ransac_mark.argtypes = [vp, vp, vp, vp, vp, i32, f64, vp, vp]
    lib.vs_tracker_cuda_stream.argtypes = [vp]
    lib.vs_tracker_cuda_stream.restype = vp
    lib.vs_tracker_set_profiling.argtypes = [vp, i32]
    lib.vs_tracker_get_profile.argtypes = [vp, i32, C.POINTER(C.c_float), C.POINTER(i32), C.POINTER(i32)]
    lib.vs_pack_image_features.argtypes = [C.POINTER(VsTrackResult), i32, C.POINTER(C.c_int32), C.POINTER(C.c_double),
                                           C.POINTER(C.c_int32)]
    lib.vs_pack_image_features.restype = C.c_int
    lib.vs_profile_kernel_name.argtypes = [i32]
    lib.vs_profile_kernel_name.restype = C.c_char_p
    pi32, pf64 = C.POINTER(i32), C.POINTER(f64)
    lib.vs_bank_create.argtypes = [i32, i32, i32, C.POINTER(vp)]
    lib.vs_bank_destroy.argtypes = [vp]
    lib.vs_bank_destroy.restype = None
    lib.vs_bank_clear.argtypes = [vp, i32]
    lib.vs_bank_add_frame.argtypes = [vp, pi32, pi32, pi32, pf64, i32, f64, f64, pi32, pi32, pf64, pi32]
    lib.vs_bank_add_frame_from_tracker.argtypes = [vp, vp, pi32, f64, f64, pi32, pi32]
    lib.vs_bank_slide.argtypes = [vp, pi32, pi32, pf64, f64]
    lib.vs_bank_triangulate.argtypes = [vp, pf64, pf64, pf64, f64, pi32]
    lib.vs_bank_get_depth_vector.argtypes = [vp, i32, i32, pf64, pi32]
    lib.vs_bank_set_depth.argtypes = [vp, i32, pf64, i32, i32]
    lib.vs_bank_get_corresponding.argtypes = [vp, i32, i32, i32, i32, pf64, pi32]
    lib.vs_bank_export.argtypes = [vp, i32, i32, pi32, pi32, pi32, pf64, pi32, pf64, pi32]
    lib.vs_bank_debug_guard_check.argtypes = [vp]
    lib.vs_bank_debug_guard_check.restype = i64
    _lib = lib
    return lib


def check(rc: int):
    if rc != VS_OK:
        raise VsError(f"vs_tracker error {rc}: {load().vs_last_error().decode(errors='replace')}")
