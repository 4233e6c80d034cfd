"""Frame ingest: the decoder must return exactly what the reference's
``cv::imread(path, cv::IMREAD_GRAYSCALE)`` (measurement_processor.cpp:185) returns, and the files -> pinned ring ->
submit/wait pipeline must publish the same ImageData records as the reference tracker fed by cv2.imread.
"""
import os
import struct
import zlib

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from mobile_slam_b200 import ingest  # noqa: E402


def _png(w, h, depth, ctype, rows, palette=None, filters=None, idat_split=1):
    """hand-assembled PNG (for the colour types / bit depths cv2.imwrite cannot produce); rows = packed scanline bytes"""
    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xffffffff)
    raw = b"".join(bytes([0 if filters is None else filters[y]]) + bytes(rows[y]) for y in range(h))
    z = zlib.compress(raw, 6)
    cut = [len(z) * i // idat_split for i in range(idat_split + 1)]
    out = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 0))
    if palette is not None:
        out += chunk(b"PLTE", bytes(np.asarray(palette, np.uint8).reshape(-1)))
    for i in range(idat_split):
        out += chunk(b"IDAT", z[cut[i]:cut[i + 1]])
    return out + chunk(b"IEND", b"")


def _filtered(rows, bpp):
    """apply PNG filters 0..4 cyclically to raw scanlines (exercises Sub / Up / Average / Paeth)"""
    out, fts, prev = [], [], np.zeros(len(rows[0]), np.int32)
    for y, r in enumerate(rows):
        cur = np.asarray(r, np.int32)
        ft = y % 5
        a = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]])
        c = np.concatenate([np.zeros(bpp, np.int32), prev[:-bpp]])
        b = prev
        if ft == 0:
            f = cur
        elif ft == 1:
            f = cur - a
        elif ft == 2:
            f = cur - b
        elif ft == 3:
            f = cur - ((a + b) >> 1)
        else:
            p = a + b - c
            pa, pb, pc = np.abs(p - a), np.abs(p - b), np.abs(p - c)
            pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, b, c))
            f = cur - pred
        out.append((f & 255).astype(np.uint8))
        fts.append(ft)
        prev = cur
    return out, fts


def _same_as_cv2(tmp_path, name, data):
    p = tmp_path / name
    p.write_bytes(data)
    want = cv2.imread(str(p), cv2.IMREAD_GRAYSCALE)
    assert want is not None, name
    got = ingest.read_gray(str(p))
    assert got.shape == want.shape and np.array_equal(got, want), f"{name}: max |d| {np.abs(got.astype(int) - want.astype(int)).max()}"


def test_png_written_by_opencv(tmp_path):
    rng = np.random.default_rng(0)
    cases = {
        "g8": rng.integers(0, 256, (37, 53), dtype=np.uint8),
        "g16": rng.integers(0, 65536, (37, 53), dtype=np.uint16),
        "rgb8": rng.integers(0, 256, (37, 53, 3), dtype=np.uint8),
        "rgba8": rng.integers(0, 256, (37, 53, 4), dtype=np.uint8),
        "rgb16": rng.integers(0, 65536, (37, 53, 3), dtype=np.uint16),
        "rgba16": rng.integers(0, 65536, (37, 53, 4), dtype=np.uint16),
        "c2_frame": rng.integers(0, 256, (480, 752), dtype=np.uint8),
    }
    for name, img in cases.items():
        for level in (1, 9):
            ok, buf = cv2.imencode(".png", img, [cv2.IMWRITE_PNG_COMPRESSION, level])
            assert ok
            _same_as_cv2(tmp_path, f"{name}_{level}.png", buf.tobytes())
    ok, buf = cv2.imencode(".png", (cases["g8"] > 128).astype(np.uint8) * 255, [cv2.IMWRITE_PNG_BILEVEL, 1])
    _same_as_cv2(tmp_path, "bilevel.png", buf.tobytes())
    ok, buf = cv2.imencode(".pgm", cases["g8"])
    _same_as_cv2(tmp_path, "p5.pgm", buf.tobytes())


def test_png_types_opencv_cannot_write(tmp_path):
    rng = np.random.default_rng(1)
    w, h = 61, 23
    # palette, 8 / 4 / 2 / 1 bits
    for depth in (8, 4, 2, 1):
        n = 1 << depth
        pal = rng.integers(0, 256, (n, 3))
        idx = rng.integers(0, n, (h, w))
        per = 8 // depth
        rows = []
        for y in range(h):
            row = np.zeros((w + per - 1) // per, np.uint8)
            for x in range(w):
                row[x // per] |= int(idx[y, x]) << (8 - depth * (1 + x % per))
            rows.append(row)
        _same_as_cv2(tmp_path, f"pal{depth}.png", _png(w, h, depth, 3, rows, palette=pal))
    # low-bit grey
    for depth in (4, 2, 1):
        n = 1 << depth
        v = rng.integers(0, n, (h, w))
        per = 8 // depth
        rows = []
        for y in range(h):
            row = np.zeros((w + per - 1) // per, np.uint8)
            for x in range(w):
                row[x // per] |= int(v[y, x]) << (8 - depth * (1 + x % per))
            rows.append(row)
        _same_as_cv2(tmp_path, f"g{depth}.png", _png(w, h, depth, 0, rows))
    # grey + alpha, 8 and 16 bits; every filter type; IDAT split over several chunks
    ga = rng.integers(0, 256, (h, w * 2), dtype=np.uint8)
    rows, fts = _filtered(list(ga), 2)
    _same_as_cv2(tmp_path, "ga8.png", _png(w, h, 8, 4, rows, filters=fts, idat_split=3))
    ga16 = rng.integers(0, 256, (h, w * 4), dtype=np.uint8)
    rows, fts = _filtered(list(ga16), 4)
    _same_as_cv2(tmp_path, "ga16.png", _png(w, h, 16, 4, rows, filters=fts))
    rgb = rng.integers(0, 256, (h, w * 3), dtype=np.uint8)
    rows, fts = _filtered(list(rgb), 3)
    _same_as_cv2(tmp_path, "rgb_filters.png", _png(w, h, 8, 2, rows, filters=fts, idat_split=4))
    g = rng.integers(0, 256, (h, w), dtype=np.uint8)
    rows, fts = _filtered(list(g), 1)
    _same_as_cv2(tmp_path, "g_filters.png", _png(w, h, 8, 0, rows, filters=fts))


def test_row_range_and_errors(tmp_path):
    from mobile_slam_b200 import VsError
    img = np.random.default_rng(2).integers(0, 256, (50, 40), dtype=np.uint8)
    p = tmp_path / "a.png"
    cv2.imwrite(str(p), img)
    assert np.array_equal(ingest.read_gray(str(p), max_h=30), img[:30])       # rowRange(0, camera.row), mp:191
    with pytest.raises(VsError, match="wide"):
        ingest.read_gray(str(p), max_w=32)
    with pytest.raises(VsError):
        ingest.decode_gray(b"not an image at all")
    bad = p.read_bytes()[:200]
    with pytest.raises(VsError):
        ingest.decode_gray(bad)
    with pytest.raises(VsError, match="interlac"):
        ingest.decode_gray(_png(4, 4, 8, 0, [np.zeros(4, np.uint8)] * 4)[:28] + b"\x01" + b"\0" * 200)


@pytest.mark.gpu
def test_files_on_disk_to_published_records(tmp_path, built_lib):
    """extractImageFeatures end to end: PNG files of 3 streams -> decode threads -> pinned ring -> submit/wait -> the
    records both callers publish, against the reference tracker reading the same files with cv2.imread."""
    from mobile_slam_b200 import BatchTracker, VS_RANSAC_OFF, config_from_synth
    from mobile_slam_b200.synth import CONFIGS, SynthStream
    from oracle import camera as ocam, cv_reference as R
    cfg = dict(CONFIGS["C2"], width=376, height=240, max_cnt=100, min_dist=15, intr=(230.0, 230.0, 188.0, 120.0))
    n_streams, n_frames = 3, 8
    W, H = cfg["width"], cfg["height"]
    paths = []
    for s in range(n_streams):
        st = SynthStream(80 + s, W, H + 16)                 # files are taller than camera.row: the tracker takes rowRange(0, row)
        row = []
        for k in range(n_frames):
            p = tmp_path / f"s{s}_{k:04d}.png"
            assert cv2.imwrite(str(p), st.frame(k))
            row.append(str(p))
        paths.append(row)
    tc, cam = config_from_synth(cfg, ransac_mode=VS_RANSAC_OFF)
    bt = BatchTracker(tc, cam, n_streams)
    pipe = ingest.IngestPipeline(bt, decode_threads=3)
    refs = [R.RefFeatureTracker(R.config_from_synth(cfg, use_ransac=False), ocam.from_config(cfg)) for _ in range(n_streams)]
    got = []
    outstanding = 0
    for k in range(n_frames):                              # keep up to three batches outstanding
        pipe.submit_files([paths[s][k] for s in range(n_streams)], [0.05 * k] * n_streams)
        outstanding += 1
        if outstanding == 3:
            got.append(pipe.wait())
            outstanding -= 1
    while outstanding:
        got.append(pipe.wait())
        outstanding -= 1
    assert len(got) == n_frames
    for k in range(n_frames):
        for s in range(n_streams):
            img = cv2.imread(paths[s][k], cv2.IMREAD_GRAYSCALE)[:H]
            refs[s].detect_and_track(img, 0.05 * k)
            refs[s].update_all_ids()
            r = got[k][s]
            assert r.ids.tolist() == refs[s].ids and r.track_cnt.tolist() == refs[s].track_cnt, (k, s)
            assert np.abs(r.cur_pts - refs[s].cur_pts).max() <= 0.01
            pub = refs[s].published()
            mine = {int(i): (float(u[0]), float(u[1]), 1.0, float(p[0]), float(p[1]), float(v[0]), float(v[1]))
                    for i, c, p, u, v in zip(r.ids, r.track_cnt, r.cur_pts, r.cur_undistorted_pts, r.pts_velocity) if c > 1}
            assert sorted(mine) == sorted(pub)
            for i in pub:
                assert np.allclose(mine[i][:5], pub[i][:5], atol=0.01) and np.allclose(mine[i][5:], pub[i][5:], atol=4e-3)
    # a missing file fails its own batch, in order, with the reference's message; the pipeline keeps working
    from mobile_slam_b200 import VsError
    pipe.submit_files([paths[s][0] for s in range(n_streams)], [1.0] * n_streams)
    pipe.submit_files([paths[0][1], str(tmp_path / "missing.png"), paths[2][1]], [1.05] * n_streams)
    pipe.submit_raw([cv2.imread(paths[s][2], cv2.IMREAD_GRAYSCALE)[:H] for s in range(n_streams)], [1.10] * n_streams)
    assert len(pipe.wait()) == n_streams
    with pytest.raises(VsError, match="Cannot load image"):
        pipe.wait()
    assert len(pipe.wait()) == n_streams
    with pytest.raises(VsError):
        pipe.wait()                                         # nothing outstanding
    pipe.close()
    bt.close()


def _fuzz_child(n_cases):
    """mutated PNG / PGM streams through vs_image_decode_gray: every call must RETURN (image or VsError)"""
    import ctypes as C
    from mobile_slam_b200 import VsError, _abi
    rng = np.random.default_rng(12345)
    img = rng.integers(0, 256, (37, 53), dtype=np.uint8)
    seeds = [cv2.imencode(".png", img)[1].tobytes(),
             cv2.imencode(".png", np.dstack([img, img[::-1], img.T[:37, :37].repeat(2, 1)[:, :53]]))[1].tobytes(),
             cv2.imencode(".png", (img.astype(np.uint16) << 8) | 7)[1].tobytes(),
             _png(9, 5, 2, 3, [rng.integers(0, 256, 3, dtype=np.uint8)] * 5, palette=rng.integers(0, 256, (4, 3))),
             cv2.imencode(".pgm", img)[1].tobytes()]
    lib = _abi.load()
    out = np.zeros((64, 64), np.uint8)
    w, h = C.c_int32(0), C.c_int32(0)
    ok = bad = 0

    def feed(data):
        nonlocal ok, bad
        buf = (C.c_ubyte * max(len(data), 1)).from_buffer_copy(data.ljust(1, b"\0"))
        rc = lib.vs_image_decode_gray(buf, len(data), out.ctypes.data, 64, 64, 64, C.byref(w), C.byref(h))
        ok, bad = ok + (rc == 0), bad + (rc != 0)

    # crafted headers first: sizes that overflow int / size_t arithmetic, zero sizes, absurd chunk lengths
    base = seeds[0]
    for wv, hv in ((0xFFFFFFFF, 1), (1, 0xFFFFFFFF), (0x7FFFFFFF, 0x7FFFFFFF), (0, 5), (5, 0), (64, 0x80000000), (0x80000001, 3)):
        hdr = struct.pack(">IIBBBBB", wv, hv, 8, 0, 0, 0, 0)
        feed(base[:16] + hdr + struct.pack(">I", zlib.crc32(b"IHDR" + hdr) & 0xffffffff) + base[33:])
    feed(base[:33] + struct.pack(">I", 0xFFFFFFF0) + base[37:])
    feed(base[:33] + struct.pack(">I", 0x7FFFFFFF) + base[37:])
    for hdr in (b"P5\n99999999999999999999 3\n255\n", b"P5\n3 99999999999\n255\n", b"P5\n2147483647 2147483647\n255\nxx",
                b"P5\n-3 4\n255\n", b"P5\n3 4\n65535\n" + b"\0" * 24, b"P5\n# only a comment", b"P5", b"P5\n3 4\n255"):
        feed(hdr)
    assert ok == 0, "a crafted header decoded"
    for k in range(n_cases):
        d = bytearray(seeds[k % len(seeds)])
        kind = rng.integers(0, 4)
        if kind == 0:                                    # byte flips
            for _ in range(int(rng.integers(1, 6))):
                d[int(rng.integers(0, len(d)))] = int(rng.integers(0, 256))
        elif kind == 1:                                  # truncation
            d = d[: int(rng.integers(0, len(d)))]
        elif kind == 2:                                  # a random 32-bit field (lengths, sizes) replaced by an extreme
            i = int(rng.integers(0, max(len(d) - 4, 1)))
            d[i:i + 4] = struct.pack(">I", int(rng.choice([0, 1, 0x7FFFFFFF, 0x80000000, 0xFFFFFFFF, 0xFFFF])))
        else:                                            # a slice duplicated / dropped
            i, j = sorted(int(x) for x in rng.integers(0, len(d), 2))
            d = d[:i] + d[j:] if rng.integers(0, 2) else d[:j] + d[i:j] + d[j:]
        feed(bytes(d))
    print(f"fuzz ok decoded={ok} rejected={bad}")


def test_decoder_survives_malformed_streams():
    """a decoder that reads files from disk must reject damage, never crash or balloon: crafted IHDR / PGM sizes that
    overflow 32-bit arithmetic, absurd chunk lengths, truncations and random mutations, in a child process"""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = f"import sys; sys.path.insert(0, {root!r}); sys.path.insert(0, {os.path.join(root, 'tests')!r}); " \
           f"import test_ingest as T; T._fuzz_child(4000)"
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root)
    assert r.returncode == 0 and "fuzz ok" in r.stdout, (r.returncode, r.stdout[-500:], r.stderr[-2000:])
