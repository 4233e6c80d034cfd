"""CPU tests of host-compilable product logic: the libstdc++ std::sort restatement the device uses to reproduce
FeatureTracker::setMask's UNSTABLE tie order (mobile_slam_b200/csrc/introsort.cuh) against the real std::sort with the
reference's comparator (feature_tracker.cpp:70-73), compiled here with g++; and the cv::circle half-width table."""
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SRC = r'''
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <utility>
#include <vector>
#include "introsort.cuh"
int main() {
    unsigned seed = 12345;
    auto rnd = [&]() { seed = seed * 1664525u + 1013904223u; return (seed >> 8); };
    int bad = 0, cases = 0;
    for (int n = 0; n <= 700; n += (n < 40 ? 1 : 37)) {
        for (int rep = 0; rep < 6; ++rep) {
            const int spread = rep < 2 ? 3 : (rep < 4 ? 12 : 1000);      // tie heavy .. tie free
            std::vector<int> key(n);
            for (int i = 0; i < n; ++i) key[i] = 1 + (int)(rnd() % spread);
            if (rep == 5) std::sort(key.begin(), key.end());              // sorted ascending: worst case for median-of-3
            std::vector<std::pair<int, int>> ref(n);
            for (int i = 0; i < n; ++i) ref[i] = {key[i], i};
            std::sort(ref.begin(), ref.end(), [](const std::pair<int, int>& a, const std::pair<int, int>& b) { return a.first > b.first; });
            std::vector<int> v(n);
            for (int i = 0; i < n; ++i) v[i] = i;
            vs::std_sort_indices(v.data(), n, vs::DescByKey{key.data()});
            ++cases;
            for (int i = 0; i < n; ++i) if (v[i] != ref[i].second) { ++bad; break; }
        }
    }
    printf("%d %d\n", cases, bad);
    return bad != 0;
}
'''


def test_introsort_restatement_matches_libstdcxx(tmp_path):
    src = tmp_path / "t.cpp"
    src.write_text(SRC)
    exe = tmp_path / "t"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-x", "c++", "-I", os.path.join(ROOT, "mobile_slam_b200", "csrc"),
                           str(src), "-o", str(exe)])
    cases, bad = (int(v) for v in subprocess.check_output([str(exe)], text=True).split())
    assert cases > 300 and bad == 0


def test_oracle_stdsort_is_the_real_std_sort():
    # the oracle's tie order comes from libstdc++ itself (oracle/stdsort.cpp), not from a guess
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
    from oracle import cv_reference as R
    tc = np.array([3, 1, 3, 2, 3, 1, 2, 2, 3, 1, 1, 1, 2, 2, 3, 3, 1, 2, 3, 1], np.int32)
    perm = R.setmask_order(tc)
    assert sorted(perm.tolist()) == list(range(len(tc)))
    assert all(tc[perm[i]] >= tc[perm[i + 1]] for i in range(len(tc) - 1))
    big = np.random.default_rng(0).integers(1, 4, 300).astype(np.int32)
    p = R.setmask_order(big)
    assert not np.array_equal(p, np.argsort(-big, kind="stable")), "std::sort is not stable on tie-heavy input"


def test_pack_image_features_matches_the_callers_loop(built_lib):
    # vs_pack_image_features = the loop of MeasurementProcessor::extractImageFeatures (measurement_processor.cpp:210-231)
    # and VIOEngine::processFrame (vio_engine.cpp:219-239): only track_cnt > 1, array order, [x, y, 1, u, v, vx, vy]
    import ctypes as C
    import numpy as np
    from mobile_slam_b200 import _abi
    lib = _abi.load()
    rng = np.random.default_rng(3)
    n = 37
    ids = np.arange(100, 100 + n, dtype=np.int32)
    cnt = rng.integers(1, 5, n).astype(np.int32)
    pts = rng.uniform(0, 700, (n, 2)).astype(np.float32)
    un = rng.uniform(-1, 1, (n, 2)).astype(np.float32)
    vel = rng.uniform(-3, 3, (n, 2)).astype(np.float32)
    r = _abi.VsTrackResult()
    r.n, r.n_new = n, 0
    r.ids = ids.ctypes.data_as(C.POINTER(C.c_int32))
    r.track_cnt = cnt.ctypes.data_as(C.POINTER(C.c_int32))
    r.cur_pts = pts.ctypes.data_as(C.POINTER(C.c_float))
    r.undist = un.ctypes.data_as(C.POINTER(C.c_float))
    r.velocity = vel.ctypes.data_as(C.POINTER(C.c_float))
    out_ids = np.zeros(n, np.int32)
    rec = np.zeros((n, 7), np.float64)
    m = C.c_int32(0)
    rc = lib.vs_pack_image_features(C.byref(r), n, out_ids.ctypes.data_as(C.POINTER(C.c_int32)),
                                    rec.ctypes.data_as(C.POINTER(C.c_double)), C.byref(m))
    assert rc == 0
    keep = cnt > 1
    assert m.value == int(keep.sum())
    assert out_ids[:m.value].tolist() == ids[keep].tolist()
    want = np.concatenate([un[keep], np.ones((m.value, 1), np.float32), pts[keep], vel[keep]], 1).astype(np.float64)
    assert np.array_equal(rec[:m.value], want)
    # too small a buffer: error code, nothing written past it
    small = np.zeros((2, 7), np.float64)
    sid = np.zeros(2, np.int32)
    rc = lib.vs_pack_image_features(C.byref(r), 2, sid.ctypes.data_as(C.POINTER(C.c_int32)),
                                    small.ctypes.data_as(C.POINTER(C.c_double)), C.byref(m))
    assert rc == _abi.VS_ERR_CAPACITY if hasattr(_abi, "VS_ERR_CAPACITY") else rc == -4
    assert m.value == 2 and np.array_equal(small, want[:2])
