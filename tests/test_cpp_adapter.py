"""The C++ host layer (include/vs_feature_tracker.hpp): compiles with g++ against the C ABI, refuses to run without CUDA
(CPU test), and on the GPU returns exactly what the Python mirror returns on the same frames (gpu test)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "mobile_slam_b200")


def _build(tmp_path, built_lib):
    exe = tmp_path / "adapter_smoke"
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "adapter_smoke.cpp"), "-o", str(exe),
                           "-L", PKG, "-lvs_tracker", f"-Wl,-rpath,{PKG}", "-lpthread"])
    return exe


def _frames(tmp_path, n=6, w=320, h=240):
    from mobile_slam_b200.synth import SynthStream
    st = SynthStream(7, w, h)
    fr = np.stack([st.frame(k) for k in range(n)])
    p = tmp_path / "frames.bin"
    fr.tofile(p)
    return p, fr


def test_adapter_compiles_and_fails_loudly_without_cuda(tmp_path, built_lib):
    import torch
    exe = _build(tmp_path, built_lib)
    p, _ = _frames(tmp_path, 2)
    r = subprocess.run([str(exe), str(p), "320", "240", "2", "-"], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stderr
    else:
        assert r.returncode == 3 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr)


def test_yaml_camera_reader_matches_python(tmp_path):
    # the adapter's FileStorage-YAML reader and mobile_slam_b200.calib agree (checked through a tiny C++ probe)
    y = tmp_path / "cam.yaml"
    y.write_text("%YAML:1.0\n---\nmodel_type: KANNALA_BRANDT\ncamera_name: cam0\nimage_width: 512\nimage_height: 512\n"
                 "projection_parameters:\n   k2: 0.0034\n   k3: 0.0007\n   k4: -0.0020\n   k5: 0.0002\n"
                 "   mu: 190.97\n   mv: 190.96\n   u0: 254.93\n   v0: 256.89\n")
    src = tmp_path / "probe.cpp"
    src.write_text('#include "vs_feature_tracker.hpp"\nint main(int, char** a){vs_camera c; if(!vs::camera_from_yaml(a[1], &c)) return 1;'
                   'std::printf("%d %d %d %.6f %.6f %.6f %.6f %.6f %.6f %.6f %.6f\\n", c.model_type, c.width, c.height, c.fx, c.fy, c.cx, c.cy,'
                   'c.d[0], c.d[1], c.d[2], c.d[3]); return 0;}\n')
    exe = tmp_path / "probe"
    subprocess.check_call(["g++", "-std=c++17", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe), "-L", PKG,
                           "-lvs_tracker", f"-Wl,-rpath,{PKG}", "-lpthread"])
    got = subprocess.check_output([str(exe), str(y)], text=True).split()
    from mobile_slam_b200.calib import camera_from_yaml
    c = camera_from_yaml(str(y))
    want = [c.model_type, c.width, c.height, c.fx, c.fy, c.cx, c.cy, *c.d]
    assert [float(v) for v in got] == pytest.approx([float(v) for v in want], abs=1e-6)


@pytest.mark.gpu
def test_adapter_matches_python_mirror(tmp_path, built_lib):
    from mobile_slam_b200 import FeatureTracker, TrackerConfig, CameraParams, VS_RANSAC_OFF, VS_CAM_PINHOLE
    exe = _build(tmp_path, built_lib)
    p, fr = _frames(tmp_path, 6)
    out = subprocess.run([str(exe), str(p), "320", "240", "6", "-"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    lines = out.stdout.strip().splitlines()
    ft = FeatureTracker(TrackerConfig(col=320, row=240, max_cnt=60, min_dist=20, ransac_mode=VS_RANSAC_OFF))
    ft.setIntrinsicParameter(VS_CAM_PINHOLE, 320, 240, 300.0, 300.0, 160.0, 120.0, -0.1, 0.01, 1e-4, -1e-4)
    for k in range(6):
        ft.detectAndTrack(fr[k], 0.05 * k)
        i = 0
        while ft.updateID(i):
            i += 1
        tok = lines[k].split()
        assert int(tok[0]) == k and int(tok[1]) == len(ft.ids)
        for j in range(min(4, len(ft.ids))):
            t = tok[2 + 6 * j: 8 + 6 * j]
            assert int(t[0]) == ft.ids[j] and int(t[5]) == ft.track_cnt[j]
            assert abs(float(t[1]) - ft.cur_pts[j][0]) < 1e-5 and abs(float(t[2]) - ft.cur_pts[j][1]) < 1e-5
            assert abs(float(t[3]) - ft.cur_undistorted_pts[j][0]) < 1e-8
    assert lines[-1].startswith("published ") and int(lines[-1].split()[1]) == len(ft.image_data())
    # stream-sharded host layer on one GPU: 5 identical streams -> identical tracks
    out = subprocess.run([str(exe), str(p), "320", "240", "6", "-", "1", "5"], capture_output=True, text=True)
    assert out.returncode == 0 and "sharded ok 1 gpus 5 streams" in out.stdout, out.stdout + out.stderr


def _build_opencv_probe(tmp_path, use_reference_headers):
    """the configuration INTEGRATION.md §2 prescribes: -DVS_WITH_OPENCV -DVS_WITH_REFERENCE_CONFIG (OpenCV = the stand-in
    of oracle/ref_shim; utility/config.h = the reference's own header where /root/reference exists, else tests/cpp/stubs)"""
    exe = tmp_path / ("probe_ref" if use_reference_headers else "probe_stub")
    inc = ["-I", "/root/reference/include"] if use_reference_headers else ["-I", os.path.join(ROOT, "tests", "cpp", "stubs")]
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-DVS_WITH_OPENCV", "-DVS_WITH_REFERENCE_CONFIG",
                           "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "oracle", "ref_shim"), *inc,
                           os.path.join(ROOT, "tests", "cpp", "adapter_opencv_probe.cpp"), "-o", str(exe),
                           "-L", PKG, "-lvs_tracker", f"-Wl,-rpath,{PKG}", "-lpthread"])
    return exe


def test_adapter_compiles_with_opencv_and_reference_config(tmp_path, built_lib):
    import torch
    exes = [_build_opencv_probe(tmp_path, False)]
    if os.path.exists("/root/reference/include/utility/config.h"):
        exes.append(_build_opencv_probe(tmp_path, True))        # against the reference's real utility/config.h
    p, _ = _frames(tmp_path, 2)
    for exe in exes:
        r = subprocess.run([str(exe), str(p), "320", "240", "2"], capture_output=True, text=True)
        if torch.cuda.is_available():
            assert r.returncode == 0, r.stderr
        else:
            assert r.returncode == 3 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr)


@pytest.mark.gpu
def test_adapter_opencv_mode_follows_g_config(tmp_path, built_lib):
    """cv::Mat frames, utility::g_config re-read on every call (max_cnt, LK criteria, equalize mutated between frames)
    == the Python mirror driven with the same changes; a frame of another size is refused."""
    from mobile_slam_b200 import FeatureTracker, TrackerConfig, VS_RANSAC_OFF, VS_CAM_PINHOLE
    exe = _build_opencv_probe(tmp_path, False)
    p, fr = _frames(tmp_path, 7)
    out = subprocess.run([str(exe), str(p), "320", "240", "7"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    lines = out.stdout.strip().splitlines()
    assert lines[-1] == "size change refused"
    ft = FeatureTracker(TrackerConfig(col=320, row=240, max_cnt=60, min_dist=20, ransac_mode=VS_RANSAC_OFF, capacity=1024))
    ft.setIntrinsicParameter(VS_CAM_PINHOLE, 320, 240, 300.0, 300.0, 160.0, 120.0, -0.1, 0.01, 1e-4, -1e-4)
    for k in range(7):
        if k == 2:
            ft.config.max_cnt = 80                     # the mirror re-reads its config on every call, like g_config
        if k == 3:
            ft.config.lk_window_size, ft.config.lk_iterations, ft.config.lk_eps = 15, 20, 0.03
        if k == 4:
            ft.config.equalize = 0
        ft.detectAndTrack(fr[k], 0.05 * k)
        i = 0
        while ft.updateID(i):
            i += 1
        tok = lines[k].split()
        assert int(tok[0]) == k and int(tok[1]) == len(ft.ids) and int(tok[2]) == len(ft.image_data()), (k, tok[:3], len(ft.ids))
        for j in range(min(4, len(ft.ids))):
            t = tok[3 + 4 * j: 7 + 4 * j]
            assert int(t[0]) == ft.ids[j] and int(t[3]) == ft.track_cnt[j]
            assert abs(float(t[1]) - ft.cur_pts[j][0]) < 1e-5 and abs(float(t[2]) - ft.cur_pts[j][1]) < 1e-5
    assert len(ft.ids) > 60          # max_cnt 80 took effect


@pytest.mark.parametrize("defs", [["-DVS_WITH_OPENCV"], ["-DVS_WITH_REFERENCE_CONFIG"]])
def test_adapter_single_macro_configurations_compile(tmp_path, defs):
    """the two mixed build modes of the adapter header (OpenCV types with the per-instance config; g_config with the
    library's own image / point types) at least type-check under -Wall -Werror"""
    src = tmp_path / "t.cpp"
    src.write_text('#include "vs_ref_callbacks.h"\n#include "utility/config.h"\n#include "vs_feature_tracker.hpp"\n'
                   'namespace utility { Config g_config; }\n'
                   'extern "C" { vs_ref_callbacks vs_ref_cb = {}; void (*vs_ref_eigvals_cb)(int, const double*, double*, double*) = nullptr; }\n'
                   'int main() { frontend::FeatureTracker ft; (void)ft; return 0; }\n')
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", *defs, "-I", os.path.join(ROOT, "include"),
                           "-I", os.path.join(ROOT, "oracle", "ref_shim"), "-I", os.path.join(ROOT, "tests", "cpp", "stubs"),
                           "-c", str(src), "-o", str(tmp_path / "t.o")])
