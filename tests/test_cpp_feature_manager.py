"""The C++ host layer of the device feature bank (include/vs_feature_manager.hpp, the reference-shaped
frontend::FeatureManager): compiles with g++ against the C ABI, with and without the Eigen / reference-config branches,
refuses to run without CUDA (CPU tests), and on the GPU takes the oracle FeatureManager's decisions on the same loop."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "mobile_slam_b200")
WIN = 10


def _build(tmp_path, built_lib):
    exe = tmp_path / "feature_manager_smoke"
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "feature_manager_smoke.cpp"), "-o", str(exe),
                           "-L", PKG, "-lvs_tracker", f"-Wl,-rpath,{PKG}", "-lpthread"])
    return exe


def test_feature_manager_adapter_compiles_and_fails_loudly_without_cuda(tmp_path, built_lib):
    import torch
    exe = _build(tmp_path, built_lib)
    r = subprocess.run([str(exe), "5", "3", "0.05"], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stderr
    else:
        assert r.returncode == 3 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr)


def test_feature_manager_adapter_eigen_and_reference_config_branches_compile(tmp_path):
    """The reference-typed overloads (VS_WITH_EIGEN) and the g_config route (VS_WITH_REFERENCE_CONFIG) type-check against
    an Eigen stand-in (oracle/ref_shim) and a stub utility/config.h; calls are instantiated, nothing runs."""
    stub = tmp_path / "utility"
    stub.mkdir()
    (stub / "config.h").write_text("namespace utility { struct E { double min_parallax = 10, init_depth = 5; }; struct C { double focal_length = 460; };"
                                   " struct Config { C camera; E estimator; }; inline Config g_config; }\n")
    src = tmp_path / "probe.cpp"
    src.write_text('#define VS_WITH_EIGEN\n#define VS_WITH_REFERENCE_CONFIG\n#include <Eigen/Dense>\n#include <map>\n#include <array>\n'
                   '#include "vs_feature_manager.hpp"\n'
                   'extern "C" { void (*vs_ref_eigvals_cb)(int, const double*, double*, double*) = nullptr; void (*vs_ref_svd_v_cb)(int, int, const double*, double*) = nullptr; }\n'
                   'struct Fr { Eigen::Matrix3d R; Eigen::Vector3d P; };\n'
                   'void use(frontend::FeatureManager& fm) {\n'
                   '  std::map<int, Eigen::Matrix<double, 7, 1>> image; fm.addFeatureAndCheckParallax(0, image);\n'
                   '  Eigen::VectorXd x(3); fm.setDepth(x); fm.clearDepth(x);\n'
                   '  std::array<Fr, 11> w; Eigen::Vector3d t; Eigen::Matrix3d r; fm.triangulateAcrossAllViews(w, t, r);\n'
                   '  fm.removeBackShiftDepth(r, t, r, t); fm.removeBack(); fm.removeFront(10); fm.removeFailures();\n'
                   '  (void)fm.getDepthVector(); (void)fm.getCorresponding(0, 1); (void)fm.getFeatureCount(); fm.clearState(); }\n'
                   'int main() { return 0; }\n')
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-fsyntax-only", "-I", str(tmp_path),
                           "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "oracle", "ref_shim"), str(src)])


def _replay(seed, n_frames, step):
    """The loop of tests/cpp/feature_manager_smoke.cpp on the oracle; returns the lines the program prints."""
    from oracle.feature_manager import FeatureManager
    state = [seed]

    def rnd():
        state[0] = (state[0] * 1103515245 + 12345) & 0x7fffffff
        return state[0]

    N = 120
    px, py, pd, gen = [], [], [], [0] * N
    for _ in range(N):
        px.append((rnd() % 4096) / 4096.0 - 0.5)
        py.append((rnd() % 4096) / 8192.0 - 0.25)
        pd.append(3.0 + (rnd() % 1024) / 128.0)
    fm = FeatureManager(10.0, 460.0, 5.0)
    eye = [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]]
    window = [(eye, [0.0, 0.0, 0.0]) for _ in range(WIN + 1)]
    frame_count, nonlinear, frames = 0, False, []
    for k in range(n_frames):
        cam_x = step * k
        image = {}
        for i in range(N):
            if rnd() % 37 == 0:
                gen[i] += 1
            x, y = px[i] - cam_x / pd[i], py[i]
            if x < -0.6 or x > 0.6:
                continue
            image[i + 1000 * gen[i]] = [x, y, 1.0, x * 460.0 + 376.0, y * 460.0 + 240.0, 0.0, 0.0]
        window[frame_count] = (eye, [cam_x, 0.0, 0.0])
        key = fm.add_feature_and_check_parallax(frame_count, image)
        frames.append(dict(k=k, fc=frame_count, key=int(key), tracked=fm.last_track_num, bank=len(fm.bank),
                           count=fm.get_feature_count(), psum=fm.last_parallax[0], pnum=fm.last_parallax[1]))
        if frame_count < WIN:
            frame_count += 1
            continue
        if nonlinear:
            fm.triangulate_across_all_views(window, [0.0, 0.0, 0.0], eye)
        if key:
            R0, P0 = window[0]
            window = window[1:] + [window[WIN]]
            if nonlinear:
                fm.remove_back_shift_depth(R0, P0, window[0][0], window[0][1])
            else:
                fm.remove_back()
        else:
            window[WIN - 1] = window[WIN]
            fm.remove_front(frame_count)
        nonlinear = True
    return fm, frames


@pytest.mark.gpu
@pytest.mark.parametrize("seed,step", [(11, 0.12), (12, 0.01)])
def test_feature_manager_adapter_equals_oracle(tmp_path, built_lib, seed, step):
    exe = _build(tmp_path, built_lib)
    n_frames = 40
    out = subprocess.run([str(exe), str(seed), str(n_frames), repr(step)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    fm, frames = _replay(seed, n_frames, step)
    lines = out.stdout.strip().splitlines()
    fl = [l.split() for l in lines if l.startswith("frame ")]
    assert len(fl) == n_frames
    for t, f in zip(fl, frames):
        got = dict(k=int(t[1]), fc=int(t[3]), key=int(t[5]), tracked=int(t[7]), bank=int(t[9]), count=int(t[11]), pnum=int(t[15]))
        want = {q: f[q] for q in got}
        assert got == want, (got, want)
        assert np.float64(float(t[13])).tobytes() == np.float64(f["psum"]).tobytes(), "parallax sum not bit-equal"
    feats = [l.split() for l in lines if l.startswith("feature ")]
    assert len(feats) == len(fm.bank) > 20
    for t, it in zip(feats, fm.bank):
        assert (int(t[1]), int(t[3]), int(t[5]), int(t[7])) == (it.feature_id, it.start_frame, len(it.obs), it.solve_flag)
        assert float(t[9]) == pytest.approx(it.estimated_depth, rel=1e-9)
        assert float(t[11]) == it.obs[0][0]
    assert any(f["key"] == 0 for f in frames[WIN:]) or step > 0.05
    assert int([l for l in lines if l.startswith("corresponding")][0].split()[1]) == len(fm.get_corresponding(WIN - 2, WIN))
    assert int([l for l in lines if l.startswith("depth_vector")][0].split()[1]) == len(fm.get_depth_vector())
