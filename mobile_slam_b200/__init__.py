"""mobile_slam_b200 — B200-native (sm_100a CUDA) VINS-Mono visual front-end.

Drop-in for the hot path of chansoopark98/Mobile-SLAM (``frontend::FeatureTracker::detectAndTrack``):
CLAHE -> LK pyramid -> pyramidal LK -> setMask -> Shi-Tomasi top-up -> liftProjective + velocity,
all as hand-written CUDA kernels behind the C ABI in ``include/vs_tracker.h``.
"""
from ._abi import (VS_RANSAC_OFF, VS_RANSAC_DEVICE, VS_RANSAC_CALLBACK, VS_CAM_KANNALA_BRANDT, VS_CAM_MEI,
                   VS_CAM_PINHOLE, VS_CAM_SCARAMUZZA, VsError, load, LIB_PATH)
from .feature_tracker import (BatchTracker, CameraParams, FeatureTracker, StreamResult, TrackerConfig,
                              clahe_no_alloc, config_from_synth)
from .feature_bank import FeatureBank, StreamFeatureManager

__all__ = ["FeatureBank", "StreamFeatureManager", "BatchTracker", "CameraParams", "FeatureTracker", "StreamResult", "TrackerConfig", "config_from_synth", "clahe_no_alloc",
           "VsError", "load", "LIB_PATH", "VS_RANSAC_OFF", "VS_RANSAC_DEVICE", "VS_RANSAC_CALLBACK",
           "VS_CAM_KANNALA_BRANDT", "VS_CAM_MEI", "VS_CAM_PINHOLE", "VS_CAM_SCARAMUZZA"]
