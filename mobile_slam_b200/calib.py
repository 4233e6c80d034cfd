"""Camera calibration files -> ``CameraParams``.

Mirrors ``CameraFactory::generateCameraFromYamlFile`` (reference src/common/camera_models/CameraFactory.cc:87-150)
and the per-model ``Parameters::readFromYamlFile`` key names (PinholeCamera.cc:103-137,
EquidistantCamera.cc:105-137, CataCamera.cc mirror_parameters/xi, ScaramuzzaCamera.cc poly/affine).
The files are OpenCV FileStorage YAML ("%YAML:1.0"); only flat ``key: value`` pairs and one level of
nesting are needed, so a small reader is used instead of a YAML library.
"""
from __future__ import annotations

from . import _abi
from .feature_tracker import CameraParams


def _read_flat_yaml(path: str) -> dict:
    out: dict = {}
    section = None
    with open(path, "r", encoding="utf-8", errors="replace") as f:
        for raw in f:
            line = raw.split("#", 1)[0].rstrip()
            if not line.strip() or line.startswith("%") or line.strip() == "---":
                continue
            indent = len(line) - len(line.lstrip())
            if ":" not in line:
                continue
            key, val = line.strip().split(":", 1)
            key, val = key.strip(), val.strip()
            if indent == 0:
                if val == "" or val.startswith("!!"):
                    section = key
                    out[section] = {}
                else:
                    section = None
                    out[key] = val.strip('"')
            elif section is not None:
                out[section][key] = val.strip('"')
    return out


def camera_from_yaml(path: str) -> CameraParams:
    y = _read_flat_yaml(path)
    model = str(y.get("model_type", "MEI")).upper()      # factory default is MEI (CameraFactory.cc:94)
    if model not in _abi.CAMERA_MODELS:
        raise ValueError(f"Unknown camera model: {model}")
    w, h = int(y.get("image_width", 0)), int(y.get("image_height", 0))
    proj = {k: float(v) for k, v in y.get("projection_parameters", {}).items()}
    dist = {k: float(v) for k, v in y.get("distortion_parameters", {}).items()}
    if model == "PINHOLE":
        return CameraParams(_abi.VS_CAM_PINHOLE, w, h, proj["fx"], proj["fy"], proj["cx"], proj["cy"],
                            (dist.get("k1", 0.0), dist.get("k2", 0.0), dist.get("p1", 0.0), dist.get("p2", 0.0)))
    if model == "KANNALA_BRANDT":
        return CameraParams(_abi.VS_CAM_KANNALA_BRANDT, w, h, proj["mu"], proj["mv"], proj["u0"], proj["v0"],
                            (proj.get("k2", 0.0), proj.get("k3", 0.0), proj.get("k4", 0.0), proj.get("k5", 0.0)))
    if model == "MEI":
        mirror = {k: float(v) for k, v in y.get("mirror_parameters", {}).items()}
        return CameraParams(_abi.VS_CAM_MEI, w, h, proj["gamma1"], proj["gamma2"], proj["u0"], proj["v0"],
                            (dist.get("k1", 0.0), dist.get("k2", 0.0), dist.get("p1", 0.0), dist.get("p2", 0.0)),
                            xi=mirror.get("xi", 0.0))
    poly = {k: float(v) for k, v in y.get("poly_parameters", {}).items()}
    aff = {k: float(v) for k, v in y.get("affine_parameters", {}).items()}
    return CameraParams(_abi.VS_CAM_SCARAMUZZA, w, h, 1.0, 1.0, aff.get("cx", 0.0), aff.get("cy", 0.0),
                        poly=tuple(poly.get(f"p{i}", 0.0) for i in range(5)),
                        affine=(aff.get("ac", 1.0), aff.get("ad", 0.0), aff.get("ae", 0.0)))
