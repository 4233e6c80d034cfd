"""CPU tests of the drop-in boundary: libvs_tracker.so loads, exports every symbol include/vs_tracker.h declares,
its structs have the layout the ctypes mirror assumes, the header is plain C, and — with no GPU — every compute
entry point fails loudly instead of falling back to a CPU path."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "vs_tracker.h")


def _declared():
    txt = open(HDR).read()
    return sorted(set(re.findall(r"VS_API\s+[^;(]*?\b(vs_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_all_exported(built_lib):
    from mobile_slam_b200 import _abi
    names = _declared()
    assert len(names) >= 26
    assert sorted(_abi.EXPORTS) == names, "ctypes EXPORTS list and the header drifted apart"
    for n in names:
        assert hasattr(built_lib, n), n
    # nothing but the C ABI is visible (kernels / helpers are hidden)
    out = subprocess.run(["nm", "-D", "--defined-only", _abi.LIB_PATH], capture_output=True, text=True).stdout
    pub = [l.split()[-1] for l in out.splitlines() if " T " in l]
    assert sorted(p for p in pub if p.startswith("vs_")) == names
    assert built_lib.vs_abi_version() == 1


def test_header_is_plain_c_and_layout_matches_ctypes(tmp_path):
    from mobile_slam_b200 import _abi
    src = tmp_path / "layout.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "vs_tracker.h"\n'
                   'int main(void){printf("%zu %zu %zu %zu %zu %zu\\n", sizeof(vs_tracker_config), sizeof(vs_camera),'
                   'sizeof(vs_track_result), offsetof(vs_tracker_config, camera), offsetof(vs_track_result, velocity),'
                   'offsetof(vs_tracker_config, fisheye_mask_image)); return 0;}\n')
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)], text=True).split()]
    assert got == [C.sizeof(_abi.VsTrackerConfig), C.sizeof(_abi.VsCamera), C.sizeof(_abi.VsTrackResult),
                   _abi.VsTrackerConfig.camera.offset, _abi.VsTrackResult.velocity.offset,
                   _abi.VsTrackerConfig.fisheye_mask_image.offset]


def test_default_config_matches_reference_defaults(built_lib):
    from mobile_slam_b200 import _abi
    c = _abi.VsTrackerConfig()
    built_lib.vs_tracker_default_config(C.byref(c))
    # include/utility/config.h:18-56 of the reference
    assert (c.max_cnt, c.min_dist, c.lk_window_size, c.lk_pyramid_levels, c.lk_iterations) == (150, 30, 21, 3, 30)
    assert (c.f_threshold, c.lk_eps, c.equalize, c.fisheye, c.focal_length) == (1.0, 0.01, 1, 0, 460.0)
    assert c.struct_size == C.sizeof(_abi.VsTrackerConfig)


def test_argument_validation_needs_no_gpu(built_lib):
    from mobile_slam_b200 import _abi
    c = _abi.VsTrackerConfig()
    built_lib.vs_tracker_default_config(C.byref(c))
    h = C.c_void_p()
    c.lk_window_size = 20                       # even window: rejected before any CUDA call
    assert built_lib.vs_tracker_create(C.byref(c), None, C.byref(h)) == -1 and not h
    assert b"lk_window_size" in built_lib.vs_last_error()
    c.lk_window_size = 21
    c.struct_size = 8
    assert built_lib.vs_tracker_create(C.byref(c), None, C.byref(h)) == -1
    assert built_lib.vs_tracker_track(None, None, 0, 0.0, None) == -1
    assert built_lib.vs_tracker_launch_count(None) == 0


def test_no_cpu_fallback_without_gpu(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from mobile_slam_b200 import BatchTracker, TrackerConfig, CameraParams, VsError
    with pytest.raises(VsError) as e:
        BatchTracker(TrackerConfig(), CameraParams(), 1)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_feature_bank_validates_arguments_and_has_no_cpu_path(built_lib):
    import torch
    h = C.c_void_p()
    assert built_lib.vs_bank_create(0, 0, 16, C.byref(h)) == -1 and not h          # VS_ERR_INVALID before any CUDA call
    assert built_lib.vs_bank_create(0, 1, 0, C.byref(h)) == -1 and not h
    assert built_lib.vs_bank_create(0, 1, 16, None) == -1
    assert built_lib.vs_bank_clear(None, 0) == -1 and b"null bank" in built_lib.vs_last_error()
    assert built_lib.vs_bank_debug_guard_check(None) == -2
    if not torch.cuda.is_available():
        assert built_lib.vs_bank_create(0, 1, 16, C.byref(h)) == -2 and not h      # VS_ERR_CUDA: no device, no fallback
        assert b"no CPU fallback" in built_lib.vs_last_error()
        from mobile_slam_b200 import FeatureBank, VsError
        with pytest.raises(VsError):
            FeatureBank(1)


def test_product_package_never_imports_oracle():
    # the product path must not route through oracle/: check statically and in a clean interpreter
    pkg = os.path.join(ROOT, "mobile_slam_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                assert "oracle" not in open(os.path.join(dp, f)).read().replace("oracle/", "").replace("the oracle", ""), f
    code = "import sys; sys.path.insert(0, %r); import mobile_slam_b200; assert not any(m == 'oracle' or m.startswith('oracle.') for m in sys.modules)" % ROOT
    subprocess.check_call([sys.executable, "-c", code])
