"""TEST INFRASTRUCTURE ONLY — numpy restatement of the OpenCV arithmetic on the hot path.

The reference (``/root/reference/src/frontend/feature_tracker.cpp``) is glue around five OpenCV
calls; OpenCV's C++ source is not vendored (only headers, ``wasm/libs/opencv``), so the algorithm
lives in a third-party dependency: OpenCV (native build unpinned "4.x", CMakeLists.txt:36; wasm
build 4.5.5).  Each function below restates one published OpenCV algorithm as SURVEY.md §8(a-spec)
S1..S6 describes it and cites the reference call site it stands in for.  ``tests/test_oracle_spec.py``
pins every function against cv2 4.13.0 (the OpenCV build in this image) on seeded inputs.

Parity status: unpinned by the reference's own tests (they hold no vectors for this path); pinned
against cv2 4.13.0 outputs instead.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def reflect101(i, n):
    """BORDER_REFLECT_101 index map, valid for -n < i < 2n-1."""
    i = np.asarray(i)
    i = np.where(i < 0, -i, i)
    return np.where(i >= n, 2 * n - 2 - i, i)


# --------------------------------------------------------------------------------------------
# S1  CLAHE — cv::CLAHE::apply, createCLAHE(3.0, Size(8,8))   feature_tracker.cpp:102-106
# --------------------------------------------------------------------------------------------
def clahe_lut(src: np.ndarray, clip_limit: float = 3.0, tiles: int = 8):
    """Per-tile LUTs, shape (tiles*tiles, 256) uint8, and the (possibly padded) source."""
    h, w = src.shape
    if h % tiles or w % tiles:
        # cv::CLAHE pads BOTH sides by tiles - (size % tiles) as soon as either one is not a multiple: a dimension that
        # already is a multiple gets a full extra `tiles` rows / columns (verified against cv2 4.13, e.g. 120 x 90)
        ph, pw = tiles - h % tiles, tiles - w % tiles
        ys = reflect101(np.arange(h + ph), h)
        xs = reflect101(np.arange(w + pw), w)
        src_e = src[np.ix_(ys, xs)]
    else:
        src_e = src
    he, we = src_e.shape
    th, tw = he // tiles, we // tiles
    area = tw * th
    clip = max(int(clip_limit * area / 256), 1)
    lut_scale = F32(255.0) / F32(area)
    luts = np.zeros((tiles * tiles, 256), np.uint8)
    for ty in range(tiles):
        for tx in range(tiles):
            tile = src_e[ty * th:(ty + 1) * th, tx * tw:(tx + 1) * tw]
            hist = np.bincount(tile.ravel(), minlength=256).astype(np.int64)
            clipped = int(np.maximum(hist - clip, 0).sum())
            hist = np.minimum(hist, clip)
            hist += clipped // 256
            res = clipped % 256
            if res:
                step = max(256 // res, 1)
                i = 0
                while i < 256 and res > 0:
                    hist[i] += 1
                    i += step
                    res -= 1
            cs = np.cumsum(hist).astype(F32)
            v = np.rint(cs * lut_scale)            # float32 product, round-half-even
            luts[ty * tiles + tx] = np.clip(v, 0, 255).astype(np.uint8)
    return luts, src_e


def clahe(src: np.ndarray, clip_limit: float = 3.0, tiles: int = 8) -> np.ndarray:
    h, w = src.shape
    luts, src_e = clahe_lut(src, clip_limit, tiles)
    he, we = src_e.shape
    th, tw = he // tiles, we // tiles
    inv_tw, inv_th = F32(1.0) / F32(tw), F32(1.0) / F32(th)

    def axis(n, inv):
        f = np.arange(n, dtype=F32) * inv - F32(0.5)
        t1 = np.floor(f).astype(np.int32)
        a = (f - t1.astype(F32)).astype(F32)
        t2 = np.minimum(t1 + 1, tiles - 1)
        t1 = np.maximum(t1, 0)
        return t1, t2, a, (F32(1.0) - a).astype(F32)

    tx1, tx2, xa, xa1 = axis(w, inv_tw)
    ty1, ty2, ya, ya1 = axis(h, inv_th)
    s = src.astype(np.intp)
    lf = luts.astype(F32)
    X1, X2 = tx1[None, :], tx2[None, :]
    Y1, Y2 = ty1[:, None], ty2[:, None]
    p1 = lf[Y1 * tiles + X1, s]
    p2 = lf[Y1 * tiles + X2, s]
    q1 = lf[Y2 * tiles + X1, s]
    q2 = lf[Y2 * tiles + X2, s]
    XA, XA1 = xa[None, :], xa1[None, :]
    YA, YA1 = ya[:, None], ya1[:, None]
    top = (p1 * XA1).astype(F32) + (p2 * XA).astype(F32)
    bot = (q1 * XA1).astype(F32) + (q2 * XA).astype(F32)
    res = (top.astype(F32) * YA1).astype(F32) + (bot.astype(F32) * YA).astype(F32)
    return np.clip(np.rint(res.astype(F32)), 0, 255).astype(np.uint8)


# --------------------------------------------------------------------------------------------
# S2  cv::buildOpticalFlowPyramid(img, pyr, win, maxLevel, withDerivatives=true)   :123
# --------------------------------------------------------------------------------------------
def pyr_down(src: np.ndarray) -> np.ndarray:
    """cv::pyrDown, 5x5 [1 4 6 4 1]^2/256, BORDER_REFLECT_101, (x + 128) >> 8."""
    h, w = src.shape
    dh, dw = (h + 1) // 2, (w + 1) // 2
    k = np.array([1, 4, 6, 4, 1], np.int32)
    s = src.astype(np.int32)
    xs = reflect101(2 * np.arange(dw)[:, None] + np.arange(-2, 3)[None, :], w)   # (dw,5)
    rows = (s[:, xs] * k[None, None, :]).sum(-1)                                   # (h,dw)
    ys = reflect101(2 * np.arange(dh)[:, None] + np.arange(-2, 3)[None, :], h)   # (dh,5)
    out = (rows[ys, :] * k[None, :, None]).sum(1)
    return ((out + 128) >> 8).astype(np.uint8)


def build_pyramid(img: np.ndarray, win: int = 21, max_level: int = 3):
    """Levels 0..L' (L' <= max_level): stops when the next level is <= win in either dimension."""
    levels = [img]
    for _ in range(max_level):
        h, w = levels[-1].shape
        nh, nw = (h + 1) // 2, (w + 1) // 2
        if nw <= win or nh <= win:
            break
        levels.append(pyr_down(levels[-1]))
    return levels


def scharr_deriv(src: np.ndarray) -> np.ndarray:
    """calcScharrDeriv: int16 (H,W,2) = (Ix, Iy), REFLECT_101 at the level's own border."""
    h, w = src.shape
    s = src.astype(np.int32)
    ym, yp = reflect101(np.arange(h) - 1, h), reflect101(np.arange(h) + 1, h)
    t0 = 3 * (s[ym] + s[yp]) + 10 * s
    t1 = s[yp] - s[ym]
    xm, xp = reflect101(np.arange(w) - 1, w), reflect101(np.arange(w) + 1, w)
    ix = t0[:, xp] - t0[:, xm]
    iy = 3 * (t1[:, xm] + t1[:, xp]) + 10 * t1
    return np.stack([ix, iy], -1).astype(np.int16)


# --------------------------------------------------------------------------------------------
# S3  cv::calcOpticalFlowPyrLK(prevPyr, nextPyr, pts, ..., win, maxLevel, COUNT|EPS)   :130-136
# --------------------------------------------------------------------------------------------
def _round_half_even_i(x):
    return np.rint(x).astype(np.int64)


def _lk_weights(a, b):
    a, b = F32(a), F32(b)
    one = F32(1.0)
    s = F32(16384.0)
    iw00 = int(np.rint(F32(F32(one - a) * F32(one - b)) * s))
    iw01 = int(np.rint(F32(a * F32(one - b)) * s))
    iw10 = int(np.rint(F32(F32(one - a) * b) * s))
    iw11 = 16384 - iw00 - iw01 - iw10
    return iw00, iw01, iw10, iw11


def lk_track(prev_levels, next_levels, pts, win: int = 21, max_count: int = 30, eps: float = 0.01,
             min_eig_threshold: float = 1e-4, return_iters: bool = False):
    """Sparse pyramidal LK.  pts (N,2) float32 -> (next_pts (N,2) f32, status (N,) u8).

    Sums are accumulated as exact integers and converted to float32 once (order independent).
    """
    pts = np.asarray(pts, F32).reshape(-1, 2)
    n = len(pts)
    nlev = min(len(prev_levels), len(next_levels))
    max_count = min(max(int(max_count), 0), 100)
    epsf = min(max(float(eps), 0.0), 10.0)
    eps2 = epsf * epsf
    hw = F32((win - 1) * 0.5)
    status = np.ones(n, np.uint8)
    nxt = np.zeros((n, 2), F32)
    iters = np.zeros((n, nlev), np.int32)
    FLT_EPS = F32(np.finfo(np.float32).eps)
    scale20 = F32(1.0 / (1 << 20))

    for level in range(nlev - 1, -1, -1):
        I = prev_levels[level]
        J = next_levels[level]
        rows, cols = I.shape
        Ip = np.pad(I.astype(np.int32), win, mode="reflect")
        Jp = np.pad(J.astype(np.int32), win, mode="reflect")
        Dp = np.pad(scharr_deriv(I).astype(np.int32), ((win, win), (win, win), (0, 0)))
        lscale = F32(1.0 / (1 << level))
        for i in range(n):
            prev_pt = (pts[i] * lscale).astype(F32)
            if level == nlev - 1:
                next_pt = prev_pt.copy()
            else:
                next_pt = (nxt[i] * F32(2.0)).astype(F32)
            nxt[i] = next_pt
            p = (prev_pt - hw).astype(F32)
            ipx, ipy = int(np.floor(p[0])), int(np.floor(p[1]))
            if ipx < -win or ipx >= cols or ipy < -win or ipy >= rows:
                if level == 0:
                    status[i] = 0
                continue
            a, b = F32(p[0] - F32(ipx)), F32(p[1] - F32(ipy))
            iw00, iw01, iw10, iw11 = _lk_weights(a, b)
            y0, x0 = ipy + win, ipx + win
            P = Ip[y0:y0 + win + 1, x0:x0 + win + 1]
            Iw = (P[:-1, :-1] * iw00 + P[:-1, 1:] * iw01 + P[1:, :-1] * iw10 + P[1:, 1:] * iw11
                  + 256) >> 9
            D = Dp[y0:y0 + win + 1, x0:x0 + win + 1]
            Dw = (D[:-1, :-1] * iw00 + D[:-1, 1:] * iw01 + D[1:, :-1] * iw10 + D[1:, 1:] * iw11
                  + 8192) >> 14
            Ix, Iy = Dw[..., 0].astype(np.int64), Dw[..., 1].astype(np.int64)
            A11 = F32(F32(int((Ix * Ix).sum())) * scale20)
            A12 = F32(F32(int((Ix * Iy).sum())) * scale20)
            A22 = F32(F32(int((Iy * Iy).sum())) * scale20)
            Dd = F32(F32(A11 * A22) - F32(A12 * A12))
            t = F32(F32(F32(A11 - A22) * F32(A11 - A22)) + F32(F32(4.0) * F32(A12 * A12)))
            min_eig = F32(F32(F32(A22 + A11) - F32(np.sqrt(t))) / F32(2 * win * win))
            if min_eig < F32(min_eig_threshold) or Dd < FLT_EPS:
                if level == 0:
                    status[i] = 0
                continue
            Dd = F32(F32(1.0) / Dd)
            next_pt = (next_pt - hw).astype(F32)
            prev_delta = np.zeros(2, F32)
            Iw64 = Iw.astype(np.int64)
            j = 0
            for j in range(max_count):
                iqx, iqy = int(np.floor(next_pt[0])), int(np.floor(next_pt[1]))
                if iqx < -win or iqx >= cols or iqy < -win or iqy >= rows:
                    if level == 0:
                        status[i] = 0
                    break
                a, b = F32(next_pt[0] - F32(iqx)), F32(next_pt[1] - F32(iqy))
                iw00, iw01, iw10, iw11 = _lk_weights(a, b)
                y0, x0 = iqy + win, iqx + win
                Q = Jp[y0:y0 + win + 1, x0:x0 + win + 1]
                Jw = (Q[:-1, :-1] * iw00 + Q[:-1, 1:] * iw01 + Q[1:, :-1] * iw10
                      + Q[1:, 1:] * iw11 + 256) >> 9
                diff = Jw.astype(np.int64) - Iw64
                b1 = F32(F32(int((diff * Ix).sum())) * scale20)
                b2 = F32(F32(int((diff * Iy).sum())) * scale20)
                dx = F32(F32(F32(A12 * b2) - F32(A22 * b1)) * Dd)
                dy = F32(F32(F32(A12 * b1) - F32(A11 * b2)) * Dd)
                delta = np.array([dx, dy], F32)
                next_pt = (next_pt + delta).astype(F32)
                nxt[i] = (next_pt + hw).astype(F32)
                iters[i, level] = j + 1
                if float(dx) * float(dx) + float(dy) * float(dy) <= eps2:   # Point2f::ddot is double
                    break
                if j > 0 and abs(F32(dx + prev_delta[0])) < 0.01 and abs(F32(dy + prev_delta[1])) < 0.01:
                    nxt[i] = (nxt[i] - delta * F32(0.5)).astype(F32)
                    break
                prev_delta = delta
    if return_iters:
        return nxt, status, iters
    return nxt, status


# --------------------------------------------------------------------------------------------
# S4  cv::goodFeaturesToTrack(img, maxCorners, 0.01, minDistance, mask)   :185-186
#     (blockSize 3, Sobel aperture 3, min-eigenvalue response)
# --------------------------------------------------------------------------------------------
def min_eig_map(img: np.ndarray, fma: bool = True) -> np.ndarray:
    """cornerMinEigenVal(blockSize=3, ksize=3), float32.

    ``fma=True`` reproduces the FMA association of cv2's SIMD body; ``False`` the plain scalar order.
    Either way the result differs from any given OpenCV build by <= ~1 ulp on a minority of pixels
     — corner *sets*, not eig bits, are the parity contract.
    """
    h, w = img.shape
    p = img.astype(np.float64)          # all products below are formed exactly, rounded to f32 once
    k1 = np.float64(F32(1.0 / 3060.0))
    k2 = np.float64(F32(2.0 / 3060.0))
    xm, xp = reflect101(np.arange(w) - 1, w), reflect101(np.arange(w) + 1, w)
    ym, yp = reflect101(np.arange(h) - 1, h), reflect101(np.arange(h) + 1, h)
    r = p[:, xp] - p[:, xm]                              # exact small ints
    if fma:
        dx = ((r[ym] + r[yp]) * k1 + (r * k2).astype(F32).astype(np.float64)).astype(F32)
        inner = (p * k2 + (p[:, xm] * k1).astype(F32).astype(np.float64)).astype(F32)
        rs = (p[:, xp] * k1 + inner.astype(np.float64)).astype(F32)
    else:
        dx = ((r * k2).astype(F32) + ((r[ym] + r[yp]) * k1).astype(F32)).astype(F32)
        rs = (((p[:, xm] * k1).astype(F32) + (p * k2).astype(F32)).astype(F32)
              + (p[:, xp] * k1).astype(F32)).astype(F32)
    dy = (rs[yp] - rs[ym]).astype(F32)
    cxx, cxy, cyy = (dx * dx).astype(F32), (dx * dy).astype(F32), (dy * dy).astype(F32)

    def box3(c):
        c = c.astype(np.float64)
        c = c[:, xm] + c + c[:, xp]
        return (c[ym] + c + c[yp]).astype(F32)

    a = box3(cxx) * F32(0.5)
    b = box3(cxy)
    c = box3(cyy) * F32(0.5)
    t = ((a - c) * (a - c)).astype(F32) + (b * b).astype(F32)
    return ((a + c).astype(F32) - np.sqrt(t.astype(F32)).astype(F32)).astype(F32)


def gftt_select(eig: np.ndarray, max_corners: int, quality: float, min_distance: float,
                mask: np.ndarray | None = None) -> np.ndarray:
    """The selection half of goodFeaturesToTrack, exact: threshold, 3x3 NMS, sort, greedy grid."""
    h, w = eig.shape
    m = np.ones((h, w), bool) if mask is None else (mask != 0)
    if not m.any():
        return np.zeros((0, 2), F32)
    max_val = eig[m].max()
    thr = F32(np.float64(max_val) * quality)
    e = np.where(eig > thr, eig, F32(0))
    pad = np.pad(e, 1, mode="constant", constant_values=-np.inf)
    dil = np.max([pad[dy:dy + h, dx:dx + w] for dy in range(3) for dx in range(3)], axis=0)
    cand = (e != 0) & (e == dil) & m
    cand[0, :] = cand[-1, :] = False
    cand[:, 0] = cand[:, -1] = False
    ys, xs = np.nonzero(cand)
    lin = ys.astype(np.int64) * w + xs
    order = np.lexsort((-lin, -e[ys, xs].astype(np.float64)))     # eig desc, address desc
    ys, xs = ys[order], xs[order]
    out = []
    if min_distance >= 1:
        cell = int(np.rint(min_distance))
        gw, gh = (w + cell - 1) // cell, (h + cell - 1) // cell
        grid = [[] for _ in range(gw * gh)]
        md2 = min_distance * min_distance
        for x, y in zip(xs.tolist(), ys.tolist()):
            cx, cy = x // cell, y // cell
            good = True
            for yy in range(max(0, cy - 1), min(gh - 1, cy + 1) + 1):
                for xx in range(max(0, cx - 1), min(gw - 1, cx + 1) + 1):
                    for (px, py) in grid[yy * gw + xx]:
                        ddx, ddy = x - px, y - py
                        if ddx * ddx + ddy * ddy < md2:
                            good = False
                            break
                    if not good:
                        break
                if not good:
                    break
            if good:
                grid[cy * gw + cx].append((x, y))
                out.append((x, y))
                if max_corners > 0 and len(out) == max_corners:
                    break
    else:
        for x, y in zip(xs.tolist(), ys.tolist()):
            out.append((x, y))
            if max_corners > 0 and len(out) == max_corners:
                break
    return np.array(out, F32).reshape(-1, 2)


def good_features_to_track(img, max_corners, quality=0.01, min_distance=30.0, mask=None, fma=True):
    return gftt_select(min_eig_map(img, fma), max_corners, quality, min_distance, mask)


# --------------------------------------------------------------------------------------------
# S5  cv::circle(mask, pt, r, value, -1)  (filled, LINE_8, shift 0)   feature_tracker.cpp:56,84
# --------------------------------------------------------------------------------------------
def circle_half_widths(radius: int) -> np.ndarray:
    """hw[|dy|] = half width of the filled-circle row at vertical offset dy (-1 = row not painted).

    Runs OpenCV's integer midpoint rasteriser once; the disc is symmetric about both axes.
    """
    r = int(radius)
    hwid = np.full(r + 1, -1, np.int32)
    dx, dy, err, plus, minus = r, 0, 0, 1, (r << 1) - 1
    while dx >= dy:
        hwid[dy] = max(hwid[dy], dx)
        hwid[dx] = max(hwid[dx], dy)
        dy += 1
        err += plus
        plus += 2
        m = -1 if err > 0 else 0           # mask = (err <= 0) - 1
        err -= minus & m
        dx += m
        minus -= m & 2
    return hwid


def fill_circle(mask: np.ndarray, cx: int, cy: int, radius: int, value: int) -> None:
    h, w = mask.shape
    hwid = circle_half_widths(radius)
    for d in range(-radius, radius + 1):
        y = cy + d
        if 0 <= y < h and hwid[abs(d)] >= 0:
            x0, x1 = max(cx - hwid[abs(d)], 0), min(cx + hwid[abs(d)], w - 1)
            if x0 <= x1:
                mask[y, x0:x1 + 1] = value


def fisheye_mask(width: int, height: int) -> np.ndarray:
    """Auto-generated circular mask, feature_tracker.cpp:50-58 (image centre, not principal point)."""
    cx, cy = width / 2.0, height / 2.0
    radius = min(cx, cy, width - cx, height - cy) * 0.95
    m = np.zeros((height, width), np.uint8)
    fill_circle(m, int(cx), int(cy), int(radius), 255)
    return m


def cv_round(x) -> int:
    """cvRound: round-half-to-even of a float/double."""
    return int(np.rint(x))
