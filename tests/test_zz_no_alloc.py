"""vs_clahe_no_alloc — the allocation-free CLAHE entry point (no counterpart in the reference,
include/vs_tracker.h) — called through the C ABI on device images and compared with the existing entry
point vs_stage_clahe (same kernels through a tracker handle) and with cv2.  Tolerance: none, CLAHE is bit-exact in every
other test of this repo (tests/test_gpu_stages.py::test_clahe_bit_exact).

Several calls in a row on one stream, with different images, image counts and sizes: the scratch (a static device array
shared by all calls) keeps the previous call's contents, which must not leak into the next result.  The last call is
bracketed by cudaMemGetInfo: an entry point that allocates nothing leaves the device's free memory where it was.

The file is named so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
torch = pytest.importorskip("torch")

from mobile_slam_b200 import BatchTracker, VS_RANSAC_OFF, config_from_synth  # noqa: E402
from mobile_slam_b200.synth import CONFIGS, SynthStream  # noqa: E402

pytestmark = pytest.mark.gpu


def _round_up(v, m):
    return (v + m - 1) // m * m


def _images(w, h, n, seed):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        if i % 3 == 0:
            img = SynthStream(seed + i, max(w, 256), max(h, 256)).frame(i)[:h, :w]
        elif i % 3 == 1:
            img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        else:                                   # low-contrast ramp: exercises clipping and the redistribution residue
            img = (np.add.outer(np.arange(h), np.arange(w)) // 7 % 40 + 100).astype(np.uint8)
        out.append(np.ascontiguousarray(img))
    return out


def _existing_entry_point(w, h, imgs):
    cfg = dict(CONFIGS["C1"], width=w, height=h, intr=(0.8 * w, 0.8 * w, w / 2, h / 2))
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    bt = BatchTracker(tc, cam, 1)
    out = [bt.stage_clahe(im) for im in imgs]
    bt.close()
    return out


def _call(lib, d_src, d_dst, w, h, n, stream):
    pitch = d_src.shape[2]
    return lib.vs_clahe_no_alloc(C.c_void_p(d_src.data_ptr()), pitch, pitch * h, C.c_void_p(d_dst.data_ptr()), pitch, pitch * h, w, h, n,
                                 C.c_void_p(stream.cuda_stream))


def test_clahe_no_alloc_equals_existing_entry_point(built_lib):
    lib = built_lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = torch.cuda.Stream()
    #         width, height, images   (the second and third calls use fewer images than the first: stale tables stay behind them)
    calls = [(752, 480, 5, 1), (333, 251, 2, 2), (333, 251, 2, 3), (512, 512, 64, 4)]
    # inputs, outputs and the expected results are all prepared BEFORE the calls, so that nothing else touches the device between them
    prepared = []
    for w, h, n, seed in calls:
        imgs = _images(w, h, n, seed)
        want = _existing_entry_point(w, h, imgs[:6])              # vs_stage_clahe (first images; all of them against cv2 below)
        pitch = _round_up(w, 16)
        host = np.full((n, h, pitch), 0xEE, np.uint8)
        for i, im in enumerate(imgs):
            host[i, :, :w] = im
        d_src = torch.from_numpy(host).to(dev)
        d_dst = torch.full((n, h, pitch), 0x55, dtype=torch.uint8, device=dev)
        prepared.append((w, h, n, imgs, want, d_src, d_dst))
    torch.cuda.synchronize()

    for k, (w, h, n, imgs, want, d_src, d_dst) in enumerate(prepared):
        last = k == len(prepared) - 2                             # the repeat of the previous shape: nothing left to load lazily
        if last:
            free_before = torch.cuda.mem_get_info()[0]
        rc = _call(lib, d_src, d_dst, w, h, n, stream)
        assert rc == 0, lib.vs_last_error().decode()
        stream.synchronize()
        if last:
            assert torch.cuda.mem_get_info()[0] == free_before, "device memory was allocated or freed during the call"
        got = d_dst.cpu().numpy()
        clahe = cv2.createCLAHE(3.0, (8, 8))
        for i in range(n):
            if i < len(want):
                assert np.array_equal(got[i, :, :w], want[i]), f"call {k} image {i}: {np.count_nonzero(got[i, :, :w] != want[i])} pixels differ from vs_stage_clahe"
            ref = clahe.apply(imgs[i])
            assert np.array_equal(got[i, :, :w], ref), f"call {k} image {i}: {np.count_nonzero(got[i, :, :w] != ref)} pixels differ from cv2"
        assert (got[:, :, w:] == 0x55).all(), "bytes between width and pitch were written"


def test_clahe_no_alloc_rejects_what_it_cannot_hold(built_lib):
    lib = built_lib
    dev = torch.device("cuda:0")
    buf = torch.zeros((2, 64, 64), dtype=torch.uint8, device=dev)
    p = buf.data_ptr()

    def call(src=p, sp=64, ss=64 * 64, dst=p + 64 * 64, dp=64, ds=64 * 64, w=64, h=64, n=1):
        rc = lib.vs_clahe_no_alloc(C.c_void_p(src), sp, ss, C.c_void_p(dst), dp, ds, w, h, n, None)
        return rc, lib.vs_last_error().decode()

    rc, msg = call(n=65)
    assert rc == -4 and "VS_CLAHE_NO_ALLOC_MAX_IMAGES" in msg and "1 .. 64" in msg          # VS_ERR_CAPACITY, nothing launched
    assert call(n=0)[0] == -4
    assert call(src=None)[0] == -1 and call(w=16)[0] == -1 and call(h=20000)[0] == -1
    assert call(sp=72)[0] == -1 and call(src=p + 4)[0] == -1 and call(dp=48)[0] == -1
    assert call(n=2, ss=64)[0] == -1
    torch.cuda.synchronize()
    assert call()[0] == 0                                                                   # and the valid call still works afterwards
    torch.cuda.synchronize()
