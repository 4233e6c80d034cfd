// K9 — sparse pyramidal Lucas-Kanade, one warp per feature, all pyramid levels in one launch.
// Replaces cv::calcOpticalFlowPyrLK at /root/reference/src/frontend/feature_tracker.cpp:130-136
// (pyramid inputs, flags = 0, minEigThreshold = 1e-4).  Arithmetic spec: SURVEY.md S3 —
// 14-bit fixed-point bilinear weights, int16 template/gradient patches, exact integer accumulation
// of the 2x2 gradient matrix and mismatch vector (order independent), float32 2x2 solve.
//
// The Scharr derivative images OpenCV precomputes are derived here from the staged u8 tile:
// derivative = 0 where the derivative sample lies outside the level (BORDER_CONSTANT), while image
// samples outside the level are REFLECT_101 — exactly the padding buildOpticalFlowPyramid applies.
#include <cstdlib>
#include <map>
#include <mutex>

#include "kernels.h"

namespace vs {

constexpr int kLkMargin = 5;          // search tile = window + 2*margin
constexpr int kLkWarpsPerCta = 4;

__host__ __device__ inline int lk_tile_size(int win) { return win + 1 + 2 * kLkMargin; }
__host__ __device__ inline int lk_tile_pitch(int win) { return (lk_tile_size(win) + 3) & ~3; }
__host__ __device__ inline int lk_warp_smem(int win) {
    return lk_tile_size(win) * lk_tile_pitch(win) + 4 * (win + 1) * (win + 1);
}

__device__ __forceinline__ long long warp_sum_exact(int v) {
    // exact 64-bit sum of 32 int32 values with two REDUX ops (16-bit halves cannot overflow)
    const int lo = v & 0xffff;
    const int hi = v >> 16;
    const int slo = __reduce_add_sync(0xffffffffu, lo);
    const int shi = __reduce_add_sync(0xffffffffu, hi);
    return ((long long)shi << 16) + (long long)slo;
}

__device__ __forceinline__ void lk_weights(float a, float b, int& w00, int& w01, int& w10, int& w11) {
    const float a1 = __fsub_rn(1.0f, a), b1 = __fsub_rn(1.0f, b);
    w00 = __float2int_rn(__fmul_rn(__fmul_rn(a1, b1), 16384.0f));
    w01 = __float2int_rn(__fmul_rn(__fmul_rn(a, b1), 16384.0f));
    w10 = __float2int_rn(__fmul_rn(__fmul_rn(a1, b), 16384.0f));
    w11 = 16384 - w00 - w01 - w10;
}

__device__ __forceinline__ void stage_tile(uint8_t* tile, int tpitch, int rows, int cols,
                                           const uint8_t* __restrict__ img, int w, int h, int pitch, int x0,
                                           int y0, int lane) {
    const bool interior = (x0 >= 0) && (y0 >= 0) && (x0 + cols <= w) && (y0 + rows <= h);
    if (interior) {
        for (int idx = lane; idx < rows * cols; idx += 32) {
            const int r = idx / cols, c = idx - r * cols;
            tile[r * tpitch + c] = __ldg(img + (int64_t)(y0 + r) * pitch + x0 + c);
        }
    } else {
        for (int idx = lane; idx < rows * cols; idx += 32) {
            const int r = idx / cols, c = idx - r * cols;
            const int gy = reflect101(y0 + r, h), gx = reflect101(x0 + c, w);
            tile[r * tpitch + c] = __ldg(img + (int64_t)gy * pitch + gx);
        }
    }
}

template <int WIN>
__global__ void __launch_bounds__(32 * kLkWarpsPerCta) lk_kernel(LkParams p) {
    pdl_entry();
    extern __shared__ __align__(16) uint8_t smem[];
    const int win = WIN ? WIN : p.win;
    constexpr int KMAX = WIN ? (WIN * WIN + 31) / 32 : (kMaxWin * kMaxWin + 31) / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int s = blockIdx.y;
    const int i = blockIdx.x * kLkWarpsPerCta + warp;
    const int n = p.counts[s];
    if (i >= n) return;

    const int ts = lk_tile_size(win), tp = lk_tile_pitch(win);
    uint8_t* tile = smem + warp * lk_warp_smem(win);
    short2* deriv = reinterpret_cast<short2*>(tile + ts * tp);
    const int dw = win + 1;
    const int npx = win * win;
    const float hw = (float)(win - 1) * 0.5f;

    const float2 pt = p.pts[(int64_t)s * p.cap + i];
    float2 next = make_float2(0.f, 0.f);
    int status = 1;
    const uint8_t* prev_base = p.prev_pyr + (int64_t)s * p.pd.stream_bytes;
    const uint8_t* next_base = p.next_pyr + (int64_t)s * p.pd.stream_bytes;

    int Iv[KMAX];        // interpolated template intensity (<= 8160)
    int Gv[KMAX];        // packed (Ix | Iy << 16), each in [-4080, 4080]

    for (int level = p.n_levels - 1; level >= 0; --level) {
        const LevelDesc L = p.pd.lv[level];
        const float lscale = 1.0f / (float)(1 << level);
        const float2 prev_pt = make_float2(__fmul_rn(pt.x, lscale), __fmul_rn(pt.y, lscale));
        if (level == p.n_levels - 1) next = prev_pt;
        else next = make_float2(__fmul_rn(next.x, 2.0f), __fmul_rn(next.y, 2.0f));

        const float ppx = __fsub_rn(prev_pt.x, hw), ppy = __fsub_rn(prev_pt.y, hw);
        const int ipx = (int)floorf(ppx), ipy = (int)floorf(ppy);
        if (ipx < -win || ipx >= L.w || ipy < -win || ipy >= L.h) {
            if (level == 0) status = 0;
            continue;
        }
        int w00, w01, w10, w11;
        lk_weights(__fsub_rn(ppx, (float)ipx), __fsub_rn(ppy, (float)ipy), w00, w01, w10, w11);

        // ---- template: stage (win+3)^2 tile at (ipx-1, ipy-1), derive Scharr, interpolate ----
        __syncwarp();
        stage_tile(tile, tp, win + 3, win + 3, prev_base + L.offset, L.w, L.h, L.pitch, ipx - 1, ipy - 1, lane);
        __syncwarp();
        for (int idx = lane; idx < dw * dw; idx += 32) {
            const int r = idx / dw, c = idx - r * dw;
            const int gx = ipx + c, gy = ipy + r;
            short2 d = make_short2(0, 0);
            if (gx >= 0 && gx < L.w && gy >= 0 && gy < L.h) {
                const uint8_t* t0 = tile + r * tp + c;
                const uint8_t* t1 = t0 + tp;
                const uint8_t* t2 = t1 + tp;
                const int a00 = t0[0], a01 = t0[1], a02 = t0[2];
                const int a10 = t1[0], a12 = t1[2];
                const int a20 = t2[0], a21 = t2[1], a22 = t2[2];
                d.x = (short)(3 * (a02 + a22 - a00 - a20) + 10 * (a12 - a10));
                d.y = (short)(3 * (a20 + a22 - a00 - a02) + 10 * (a21 - a01));
            }
            deriv[idx] = d;
        }
        __syncwarp();
        int s11 = 0, s12 = 0, s22 = 0;
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
            const int q = lane + 32 * k;
            Iv[k] = 0;
            Gv[k] = 0;
            if (q < npx) {
                const int y = q / win, x = q - y * win;
                const uint8_t* t = tile + (y + 1) * tp + (x + 1);
                const int iv = (t[0] * w00 + t[1] * w01 + t[tp] * w10 + t[tp + 1] * w11 + 256) >> 9;
                const short2 d00 = deriv[y * dw + x], d01 = deriv[y * dw + x + 1];
                const short2 d10 = deriv[(y + 1) * dw + x], d11 = deriv[(y + 1) * dw + x + 1];
                const int ix = (d00.x * w00 + d01.x * w01 + d10.x * w10 + d11.x * w11 + 8192) >> 14;
                const int iy = (d00.y * w00 + d01.y * w01 + d10.y * w10 + d11.y * w11 + 8192) >> 14;
                Iv[k] = iv;
                Gv[k] = (ix & 0xffff) | (iy << 16);
                s11 += ix * ix;
                s12 += ix * iy;
                s22 += iy * iy;
            }
        }
        const float kScale = 1.0f / (float)(1 << 20);
        const float A11 = __fmul_rn(__ll2float_rn(warp_sum_exact(s11)), kScale);
        const float A12 = __fmul_rn(__ll2float_rn(warp_sum_exact(s12)), kScale);
        const float A22 = __fmul_rn(__ll2float_rn(warp_sum_exact(s22)), kScale);
        float D = __fsub_rn(__fmul_rn(A11, A22), __fmul_rn(A12, A12));
        const float dA = __fsub_rn(A11, A22);
        const float disc = __fadd_rn(__fmul_rn(dA, dA), __fmul_rn(__fmul_rn(4.0f, A12), A12));
        const float min_eig = __fdiv_rn(__fsub_rn(__fadd_rn(A22, A11), __fsqrt_rn(disc)), (float)(2 * win * win));
        if (min_eig < p.min_eig_threshold || D < 1.1920928955078125e-07f) {
            if (level == 0) status = 0;
            continue;
        }
        D = __fdiv_rn(1.0f, D);

        // ---- iterations on the next image ----
        float qx = __fsub_rn(next.x, hw), qy = __fsub_rn(next.y, hw);
        float pdx = 0.f, pdy = 0.f;
        const uint8_t* J = next_base + L.offset;
        int tox = 0, toy = 0;
        bool staged = false;
        for (int j = 0; j < p.max_count; ++j) {
            const int iqx = (int)floorf(qx), iqy = (int)floorf(qy);
            if (iqx < -win || iqx >= L.w || iqy < -win || iqy >= L.h) {
                if (level == 0) status = 0;
                break;
            }
            if (!staged || iqx < tox || iqx > tox + 2 * kLkMargin || iqy < toy || iqy > toy + 2 * kLkMargin) {
                tox = iqx - kLkMargin;
                toy = iqy - kLkMargin;
                __syncwarp();
                stage_tile(tile, tp, ts, ts, J, L.w, L.h, L.pitch, tox, toy, lane);
                __syncwarp();
                staged = true;
            }
            lk_weights(__fsub_rn(qx, (float)iqx), __fsub_rn(qy, (float)iqy), w00, w01, w10, w11);
            const uint8_t* tb = tile + (iqy - toy) * tp + (iqx - tox);
            int sb1 = 0, sb2 = 0;
#pragma unroll
            for (int k = 0; k < KMAX; ++k) {
                const int q = lane + 32 * k;
                if (q < npx) {
                    const int y = q / win, x = q - y * win;
                    const uint8_t* t = tb + y * tp + x;
                    const int jv = (t[0] * w00 + t[1] * w01 + t[tp] * w10 + t[tp + 1] * w11 + 256) >> 9;
                    const int diff = jv - Iv[k];
                    const int ix = (int)(short)(Gv[k] & 0xffff), iy = Gv[k] >> 16;
                    sb1 += diff * ix;
                    sb2 += diff * iy;
                }
            }
            const float b1 = __fmul_rn(__ll2float_rn(warp_sum_exact(sb1)), kScale);
            const float b2 = __fmul_rn(__ll2float_rn(warp_sum_exact(sb2)), kScale);
            const float dx = __fmul_rn(__fsub_rn(__fmul_rn(A12, b2), __fmul_rn(A22, b1)), D);
            const float dy = __fmul_rn(__fsub_rn(__fmul_rn(A12, b1), __fmul_rn(A11, b2)), D);
            qx = __fadd_rn(qx, dx);
            qy = __fadd_rn(qy, dy);
            next = make_float2(__fadd_rn(qx, hw), __fadd_rn(qy, hw));
            if ((double)dx * (double)dx + (double)dy * (double)dy <= p.eps2) break;
            if (j > 0 && fabs((double)__fadd_rn(dx, pdx)) < 0.01 && fabs((double)__fadd_rn(dy, pdy)) < 0.01) {
                next.x = __fsub_rn(next.x, __fmul_rn(dx, 0.5f));
                next.y = __fsub_rn(next.y, __fmul_rn(dy, 0.5f));
                break;
            }
            pdx = dx;
            pdy = dy;
        }
    }
    if (lane == 0) {
        p.next_pts[(int64_t)s * p.cap + i] = next;
        p.status[(int64_t)s * p.cap + i] = (uint8_t)status;
    }
}

// =================================================================================================
// Fast path (win = 21): every lane owns two horizontal runs of 7 window pixels (21 rows x 3 runs = 63
// runs for 32 lanes x 2).  Tiles are staged as 16-byte chunks (cp.async); a run reads 3 words per tile
// row and re-aligns them with funnel shifts, the bilinear blend is two DP2A per pixel (16-bit
// weights x u8 pixels), the Scharr derivatives five DP4A per sample, and the mismatch vector is
// accumulated as sum(J*Ix) - sum(I*Ix) with the second term fixed per level.  All integer and
// exact, so results are bit-identical to the generic kernel above.
// =================================================================================================
constexpr int kFastPitch = 48;                 // tile pitch in bytes: three 16-byte chunks per row
constexpr int kFastRows = 32;                  // search tile rows = win + 1 + 2 * margin
constexpr int kFastWarpSmem = kFastPitch * kFastRows;

__device__ __forceinline__ int dp2a_lo_su(int a, unsigned b, int c) {
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_hi_su(int a, unsigned b, int c) {
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp4a_us(unsigned a, int b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// Stage NROWS rows of 48 bytes of a level into the warp's tile; x0 is a multiple of 16.  A 16-byte chunk goes global ->
// shared with ONE asynchronous copy (cp.async.cg, 128 bits, no register staging): 3 instructions per lane for a whole
// search tile.  Levels >= kPyrFirstPaddedLevel carry a REFLECT_101 border (kPyrPadX x kPyrPadY, pyr_pad_*_kernel) that
// covers every tile this kernel can ask for (origins down to -48 columns / -26 rows, ends up to w + 42 / h + 26), so
// there PADDED = true and no chunk needs border handling.  On the unpadded level 0 a chunk that is not fully inside the
// level is assembled from REFLECT_101 bytes.  Pitches and level offsets are multiples of 16: chunks are aligned on both sides.
template <int NROWS, bool PADDED>
__device__ __forceinline__ void stage_rows16(uint32_t* tile, const uint8_t* __restrict__ img, int w, int h, int pitch,
                                             int x0, int y0, int lane) {
    constexpr int CH = kFastPitch / 16, TOTAL = NROWS * CH, U = (TOTAL + 31) / 32;
    static_assert(kPyrPadX >= kFastPitch && kPyrPadY >= 21 + 1 + kLkMargin, "level border too small for the tiles");
    const uint32_t tile_s = (uint32_t)__cvta_generic_to_shared(tile);
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const int idx = lane + 32 * u;
        if (U * 32 != TOTAL && idx >= TOTAL) break;
        const int r = idx / CH, c = idx - r * CH;
        const int gx = x0 + 16 * c, gy = y0 + r;
        if (PADDED || (gy >= 0 && gy < h && gx >= 0 && gx + 16 <= w)) {
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(tile_s + (uint32_t)(r * kFastPitch + 16 * c)),
                         "l"(img + (int64_t)gy * pitch + gx) : "memory");
        } else {
            const uint8_t* row = img + (int64_t)reflect101(gy, h) * pitch;
            uint32_t* dst = tile + r * (kFastPitch / 4) + 4 * c;
#pragma unroll 1
            for (int q = 0; q < 4; ++q) {
                const int x = gx + 4 * q;
                uint32_t v;
                if (x >= 0 && x + 3 < w) {
                    v = __ldg(reinterpret_cast<const uint32_t*>(row + x));
                } else {
                    v = (uint32_t)__ldg(row + reflect101(x, w)) | ((uint32_t)__ldg(row + reflect101(x + 1, w)) << 8) |
                        ((uint32_t)__ldg(row + reflect101(x + 2, w)) << 16) | ((uint32_t)__ldg(row + reflect101(x + 3, w)) << 24);
                }
                dst[q] = v;
            }
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
}

// The same tile by ONE bulk tensor copy (TMA): lane 0 arms the warp's mbarrier with the byte count and issues
// cp.async.bulk.tensor.3d (x = byte column, y = row, z = stream) on the tensor map of the level's padded block; every lane
// then waits for the barrier's phase.  ROWS selects the map (24 template rows / 32 search rows).  A/B variant of
// stage_rows16 (VS_LK_TMA=1): same bytes in the same layout, one instruction instead of three per lane.
struct TmaWarp {
    uint32_t bar;       // shared address of this warp's mbarrier
    uint32_t phase;
};
template <int ROWS>
__device__ __forceinline__ void stage_rows_tma(uint32_t* tile, const CUtensorMap* map, int x0, int y0, int stream, int lane,
                                               TmaWarp& tw) {
    const uint32_t tile_s = (uint32_t)__cvta_generic_to_shared(tile);
    if (lane == 0) {
        // the tile was last touched through the generic proxy (ld.shared of the previous iteration): order those before
        // the async-proxy write
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tw.bar), "r"(ROWS * kFastPitch) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(tile_s), "l"(map), "r"(x0 + kPyrPadX), "r"(y0 + kPyrPadY), "r"(stream), "r"(tw.bar) : "memory");
    }
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(tw.bar), "r"(tw.phase) : "memory");
    }
    tw.phase ^= 1u;
}

// 12 bytes of a tile row starting at byte `sb`, re-aligned so that a0 holds bytes sb..sb+3, a1 sb+4..sb+7, a2 sb+8..sb+11
__device__ __forceinline__ void load_aligned3(const uint32_t* trow, int sb, uint32_t& a0, uint32_t& a1, uint32_t& a2) {
    const uint32_t* q = trow + (sb >> 2);
    const int sh = (sb & 3) * 8;
    const uint32_t w0 = q[0], w1 = q[1], w2 = q[2];
    a0 = __funnelshift_r(w0, w1, sh);
    a1 = __funnelshift_r(w1, w2, sh);
    a2 = w2 >> sh;
}

// one feature (stream s, slot i) on one warp
template <bool TMA>
__device__ __forceinline__ void lk_fast21_feature(const LkParams& p, int s, int i, uint32_t* tile, int lane,
                                                  const LkTmaMaps* maps, TmaWarp& tw) {
    constexpr int WIN = 21, NSEG = 3, NRUN = WIN * NSEG;
    const float hw = (float)(WIN - 1) * 0.5f;

    // this lane's two runs: run r -> window row r / 3, first window column 7 * (r % 3)
    int rrow[2], rcol[2];
    bool rok[2];
#pragma unroll
    for (int t = 0; t < 2; ++t) {
        const int r = lane + 32 * t;
        rok[t] = r < NRUN;
        const int rr = rok[t] ? r : 0;
        rrow[t] = rr / NSEG;
        rcol[t] = 7 * (rr - rrow[t] * NSEG);
    }

    const float2 pt = p.pts[(int64_t)s * p.cap + i];
    float2 next = make_float2(0.f, 0.f);
    int status = 1;
    const uint8_t* prev_base = p.prev_pyr + (int64_t)s * p.pd.stream_bytes;
    const uint8_t* next_base = p.next_pyr + (int64_t)s * p.pd.stream_bytes;
    const int C3 = 0x000300FD, C10 = 0x000A00F6;        // Scharr x: (-3,0,3), (-10,0,10) as s8x4
    const int CYP = 0x00030A03, CYN = 0x00FDF6FD;       // Scharr y: +(3,10,3) on row+1, -(3,10,3) on row-1

    int ix[2][7], iy[2][7];      // interpolated template gradients
    int c1[2], c2[2];            // sum(I * Ix), sum(I * Iy) of each run

    for (int level = p.n_levels - 1; level >= 0; --level) {
        const LevelDesc L = p.pd.lv[level];
        const float lscale = 1.0f / (float)(1 << level);
        const float2 prev_pt = make_float2(__fmul_rn(pt.x, lscale), __fmul_rn(pt.y, lscale));
        if (level == p.n_levels - 1) next = prev_pt;
        else next = make_float2(__fmul_rn(next.x, 2.0f), __fmul_rn(next.y, 2.0f));

        const float ppx = __fsub_rn(prev_pt.x, hw), ppy = __fsub_rn(prev_pt.y, hw);
        const int ipx = (int)floorf(ppx), ipy = (int)floorf(ppy);
        if (ipx < -WIN || ipx >= L.w || ipy < -WIN || ipy >= L.h) {
            if (level == 0) status = 0;
            continue;
        }
        int w00, w01, w10, w11;
        lk_weights(__fsub_rn(ppx, (float)ipx), __fsub_rn(ppy, (float)ipy), w00, w01, w10, w11);
        int wt = (w00 & 0xffff) | (w01 << 16), wb = (w10 & 0xffff) | (w11 << 16);

        // ---- template: tile origin (x aligned down to 16, ipy - 1), 24 rows x 48 bytes ----
        const int tx0 = (ipx - 1) & ~15, ax = (ipx - 1) - tx0;
        __syncwarp();
        if constexpr (TMA) stage_rows_tma<WIN + 3>(tile, &maps->m[p.prev_buf][level][0], tx0, ipy - 1, s, lane, tw);
        else if (level >= kPyrFirstPaddedLevel) stage_rows16<WIN + 3, true>(tile, prev_base + L.offset, L.w, L.h, L.pitch, tx0, ipy - 1, lane);
        else stage_rows16<WIN + 3, false>(tile, prev_base + L.offset, L.w, L.h, L.pitch, tx0, ipy - 1, lane);
        __syncwarp();
        const bool border = ipx < 0 || ipy < 0 || ipx + WIN + 1 > L.w || ipy + WIN + 1 > L.h;
        int s11 = 0, s12 = 0, s22 = 0;
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            c1[t] = 0;
            c2[t] = 0;
            // derivative samples j = 0..7 of window rows rrow, rrow+1 need tile rows rrow .. rrow+3 and
            // bytes sb .. sb+9 with sb = ax + rcol (one left of the first sample)
            const int sb = ax + rcol[t];
            uint32_t S[4][8];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                // 10 bytes sb .. sb+9 can straddle four aligned words
                const uint32_t* q = tile + (rrow[t] + r) * (kFastPitch / 4) + (sb >> 2);
                const int sh = (sb & 3) * 8;
                const uint32_t a0 = __funnelshift_r(q[0], q[1], sh), a1 = __funnelshift_r(q[1], q[2], sh),
                               a2 = __funnelshift_r(q[2], q[3], sh);
                S[r][0] = a0;
                S[r][1] = __funnelshift_r(a0, a1, 8);
                S[r][2] = __funnelshift_r(a0, a1, 16);
                S[r][3] = __funnelshift_r(a0, a1, 24);
                S[r][4] = a1;
                S[r][5] = __funnelshift_r(a1, a2, 8);
                S[r][6] = __funnelshift_r(a1, a2, 16);
                S[r][7] = __funnelshift_r(a1, a2, 24);
            }
            int dx0[8], dy0[8], dx1[8], dy1[8];   // derivative rows rrow and rrow + 1
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                dx0[j] = dp4a_us(S[2][j], C3, dp4a_us(S[1][j], C10, dp4a_us(S[0][j], C3, 0)));
                dx1[j] = dp4a_us(S[3][j], C3, dp4a_us(S[2][j], C10, dp4a_us(S[1][j], C3, 0)));
                dy0[j] = dp4a_us(S[2][j], CYP, dp4a_us(S[0][j], CYN, 0));
                dy1[j] = dp4a_us(S[3][j], CYP, dp4a_us(S[1][j], CYN, 0));
            }
            if (border) {   // derivative image is zero-padded (BORDER_CONSTANT) outside the level
                const int gy0 = ipy + rrow[t], gx0 = ipx + rcol[t];
                const bool y0ok = gy0 >= 0 && gy0 < L.h, y1ok = gy0 + 1 >= 0 && gy0 + 1 < L.h;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const bool xok = gx0 + j >= 0 && gx0 + j < L.w;
                    if (!(xok && y0ok)) { dx0[j] = 0; dy0[j] = 0; }
                    if (!(xok && y1ok)) { dx1[j] = 0; dy1[j] = 0; }
                }
            }
#pragma unroll
            for (int j = 0; j < 7; ++j) {
                const int gx = (dx0[j] * w00 + dx0[j + 1] * w01 + dx1[j] * w10 + dx1[j + 1] * w11 + 8192) >> 14;
                const int gy = (dy0[j] * w00 + dy0[j + 1] * w01 + dy1[j] * w10 + dy1[j + 1] * w11 + 8192) >> 14;
                // template intensity: bytes (x, x+1) of tile rows rrow+1, rrow+2 = low half of S[.][j+1]
                const int iv = dp2a_lo_su(wb, S[2][j + 1], dp2a_lo_su(wt, S[1][j + 1], 256)) >> 9;
                const bool ok = rok[t];
                ix[t][j] = ok ? gx : 0;
                iy[t][j] = ok ? gy : 0;
                s11 += ix[t][j] * ix[t][j];
                s12 += ix[t][j] * iy[t][j];
                s22 += iy[t][j] * iy[t][j];
                c1[t] += iv * ix[t][j];
                c2[t] += iv * iy[t][j];
            }
        }
        const float kScale = 1.0f / (float)(1 << 20);
        const float A11 = __fmul_rn(__ll2float_rn(warp_sum_exact(s11)), kScale);
        const float A12 = __fmul_rn(__ll2float_rn(warp_sum_exact(s12)), kScale);
        const float A22 = __fmul_rn(__ll2float_rn(warp_sum_exact(s22)), kScale);
        float D = __fsub_rn(__fmul_rn(A11, A22), __fmul_rn(A12, A12));
        const float dA = __fsub_rn(A11, A22);
        const float disc = __fadd_rn(__fmul_rn(dA, dA), __fmul_rn(__fmul_rn(4.0f, A12), A12));
        const float min_eig = __fdiv_rn(__fsub_rn(__fadd_rn(A22, A11), __fsqrt_rn(disc)), (float)(2 * WIN * WIN));
        if (min_eig < p.min_eig_threshold || D < 1.1920928955078125e-07f) {
            if (level == 0) status = 0;
            continue;
        }
        D = __fdiv_rn(1.0f, D);

        // ---- iterations on the next image ----
        float qx = __fsub_rn(next.x, hw), qy = __fsub_rn(next.y, hw);
        float pdx = 0.f, pdy = 0.f;
        const uint8_t* J = next_base + L.offset;
        int tox = 0, toy = 0;          // image coordinates of tile byte (0, 0); tox is a multiple of 16
        bool staged = false;
        for (int j = 0; j < p.max_count; ++j) {
            const int iqx = (int)floorf(qx), iqy = (int)floorf(qy);
            if (iqx < -WIN || iqx >= L.w || iqy < -WIN || iqy >= L.h) {
                if (level == 0) status = 0;
                break;
            }
            // the tile must hold columns iqx .. iqx+21 (+ alignment slack) and rows iqy .. iqy+21
            if (!staged || iqx < tox || iqx + WIN + 1 + 3 > tox + kFastPitch || iqy < toy || iqy + WIN + 1 > toy + kFastRows) {
                tox = (iqx - kLkMargin) & ~15;
                toy = iqy - kLkMargin;
                __syncwarp();
                if constexpr (TMA) stage_rows_tma<kFastRows>(tile, &maps->m[p.next_buf][level][1], tox, toy, s, lane, tw);
                else if (level >= kPyrFirstPaddedLevel) stage_rows16<kFastRows, true>(tile, J, L.w, L.h, L.pitch, tox, toy, lane);
                else stage_rows16<kFastRows, false>(tile, J, L.w, L.h, L.pitch, tox, toy, lane);
                __syncwarp();
                staged = true;
            }
            lk_weights(__fsub_rn(qx, (float)iqx), __fsub_rn(qy, (float)iqy), w00, w01, w10, w11);
            wt = (w00 & 0xffff) | (w01 << 16);
            wb = (w10 & 0xffff) | (w11 << 16);
            int sb1 = 0, sb2 = 0;
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const int sb = iqx - tox + rcol[t];
                const uint32_t* trow = tile + (iqy - toy + rrow[t]) * (kFastPitch / 4);
                uint32_t a0, a1, a2, b0, b1, b2;
                load_aligned3(trow, sb, a0, a1, a2);
                load_aligned3(trow + kFastPitch / 4, sb, b0, b1, b2);
                const uint32_t a01 = __funnelshift_r(a0, a1, 8), a12 = a1 >> 8;
                const uint32_t b01 = __funnelshift_r(b0, b1, 8), b12 = b1 >> 8;
                int jv[7];
                jv[0] = dp2a_lo_su(wb, b0, dp2a_lo_su(wt, a0, 256));
                jv[1] = dp2a_lo_su(wb, b01, dp2a_lo_su(wt, a01, 256));
                jv[2] = dp2a_hi_su(wb, b0, dp2a_hi_su(wt, a0, 256));
                jv[3] = dp2a_hi_su(wb, b01, dp2a_hi_su(wt, a01, 256));
                jv[4] = dp2a_lo_su(wb, b1, dp2a_lo_su(wt, a1, 256));
                jv[5] = dp2a_lo_su(wb, b12, dp2a_lo_su(wt, a12, 256));
                jv[6] = dp2a_hi_su(wb, b1, dp2a_hi_su(wt, a1, 256));
                int t1 = -c1[t], t2 = -c2[t];
#pragma unroll
                for (int k = 0; k < 7; ++k) {
                    const int v = jv[k] >> 9;
                    t1 += v * ix[t][k];
                    t2 += v * iy[t][k];
                }
                sb1 += t1;
                sb2 += t2;
            }
            const float b1 = __fmul_rn(__ll2float_rn(warp_sum_exact(sb1)), kScale);
            const float b2 = __fmul_rn(__ll2float_rn(warp_sum_exact(sb2)), kScale);
            const float dx = __fmul_rn(__fsub_rn(__fmul_rn(A12, b2), __fmul_rn(A22, b1)), D);
            const float dy = __fmul_rn(__fsub_rn(__fmul_rn(A12, b1), __fmul_rn(A11, b2)), D);
            qx = __fadd_rn(qx, dx);
            qy = __fadd_rn(qy, dy);
            next = make_float2(__fadd_rn(qx, hw), __fadd_rn(qy, hw));
            if ((double)dx * (double)dx + (double)dy * (double)dy <= p.eps2) break;
            if (j > 0 && fabs((double)__fadd_rn(dx, pdx)) < 0.01 && fabs((double)__fadd_rn(dy, pdy)) < 0.01) {
                next.x = __fsub_rn(next.x, __fmul_rn(dx, 0.5f));
                next.y = __fsub_rn(next.y, __fmul_rn(dy, 0.5f));
                break;
            }
            pdx = dx;
            pdy = dy;
        }
    }
    if (lane == 0) {
        p.next_pts[(int64_t)s * p.cap + i] = next;
        p.status[(int64_t)s * p.cap + i] = (uint8_t)status;
    }
}

// Persistent warps: features finish after very different iteration counts (1 .. 30 per level), so a fixed
// feature -> warp map leaves CTAs waiting for their slowest warp.  Every warp instead pulls the next
// (stream, slot) work item from a global counter until the batch is drained.  Two counters alternate between
// launches: launch k drains counter k&1 and clears the other one for launch k+1 (stream order makes that safe).
__global__ void __launch_bounds__(32 * kLkWarpsPerCta, 5) lk_fast21_kernel(LkParams p, int n_streams, int max_features,
                                                                        int* counters, int which) {
    pdl_entry();
    __shared__ __align__(16) uint32_t s_tiles[kLkWarpsPerCta][kFastWarpSmem / 4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (blockIdx.x == 0 && threadIdx.x == 0) counters[which ^ 1] = 0;
    uint32_t* tile = s_tiles[warp];
    const int total = n_streams * max_features;
    int* ctr = counters + which;
    for (;;) {
        int w = 0;
        if (lane == 0) w = atomicAdd(ctr, 1);
        w = __shfl_sync(0xffffffffu, w, 0);
        if (w >= total) break;
        const int s = w / max_features, i = w - s * max_features;
        TmaWarp none{0u, 0u};
        if (i < __ldg(p.counts + s)) lk_fast21_feature<false>(p, s, i, tile, lane, nullptr, none);
        __syncwarp();
    }
}

// The same kernel with the tiles staged by TMA (VS_LK_TMA=1): the tensor maps travel as a __grid_constant__ parameter
// (3 KB), each warp owns one mbarrier.
__global__ void __launch_bounds__(32 * kLkWarpsPerCta, 5) lk_fast21_tma_kernel(LkParams p, int n_streams, int max_features,
                                                                            int* counters, int which,
                                                                            const __grid_constant__ LkTmaMaps maps) {
    pdl_entry();
    __shared__ __align__(128) uint32_t s_tiles[kLkWarpsPerCta][kFastWarpSmem / 4];
    __shared__ __align__(8) unsigned long long s_bar[kLkWarpsPerCta];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (blockIdx.x == 0 && threadIdx.x == 0) counters[which ^ 1] = 0;
    uint32_t* tile = s_tiles[warp];
    TmaWarp tw{(uint32_t)__cvta_generic_to_shared(&s_bar[warp]), 0u};
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tw.bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const int total = n_streams * max_features;
    int* ctr = counters + which;
    for (;;) {
        int w = 0;
        if (lane == 0) w = atomicAdd(ctr, 1);
        w = __shfl_sync(0xffffffffu, w, 0);
        if (w >= total) break;
        const int s = w / max_features, i = w - s * max_features;
        if (i < __ldg(p.counts + s)) lk_fast21_feature<true>(p, s, i, tile, lane, &maps, tw);
        __syncwarp();
    }
}

int lk_build_tma_maps(LkTmaMaps* out, uint8_t* const d_pyr[2], const PyrDesc& pd, int n_streams) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn) return -1;
    EncodeFn encode = reinterpret_cast<EncodeFn>(fn);
    static_assert(kPyrFirstPaddedLevel == 0, "the TMA variant addresses every level through its padded block");
    for (int b = 0; b < 2; ++b)
        for (int l = 0; l < kMaxLevels; ++l)
            for (int k = 0; k < 2; ++k) {
                const LevelDesc& L = pd.lv[l];
                void* base = d_pyr[b] + (L.offset - (int64_t)kPyrPadY * L.pitch - kPyrPadX);
                const cuuint64_t dims[3] = {(cuuint64_t)L.pitch, (cuuint64_t)(L.h + 2 * kPyrPadY), (cuuint64_t)n_streams};
                const cuuint64_t strides[2] = {(cuuint64_t)L.pitch, (cuuint64_t)pd.stream_bytes};
                const cuuint32_t box[3] = {(cuuint32_t)kFastPitch, (cuuint32_t)(k == 0 ? 21 + 3 : kFastRows), 1u};
                const cuuint32_t estr[3] = {1u, 1u, 1u};
                if (encode(&out->m[b][l][k], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
                    return -2;
            }
    return 0;
}

int launch_lk(const LkParams& p, int n_streams, int max_features, cudaStream_t st, int64_t* launches) {
    if (p.win < 3 || p.win > kMaxWin || (p.win & 1) == 0) return -1;
    if (max_features <= 0 || p.n_levels <= 0) return 0;
    dim3 grid((max_features + kLkWarpsPerCta - 1) / kLkWarpsPerCta, n_streams);
    if (p.win == 21 && p.work_counters) {
        // resident CTAs of the persistent kernel, per device (one process may drive several GPUs: vs::ShardedTracker)
        static std::mutex mu;
        static std::map<int, int> resident_of;
        int dev = 0;
        cudaGetDevice(&dev);
        int resident;
        {
            std::lock_guard<std::mutex> lock(mu);
            int& r = resident_of[dev];
            if (!r) {
                int sms = 148, per_sm = 2;
                cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
                cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lk_fast21_kernel, 32 * kLkWarpsPerCta, 0);
                r = sms * (per_sm > 0 ? per_sm : 1);
            }
            resident = r;
        }
        const int need = (n_streams * max_features + kLkWarpsPerCta - 1) / kLkWarpsPerCta;
        // VS_LK_TMA=1: tiles staged by cp.async.bulk.tensor instead of cp.async (A/B)
        static const bool want_tma = [] { const char* e = getenv("VS_LK_TMA"); return e && atoi(e) != 0; }();
        if (want_tma && p.tma_maps)
            launch_pdl(lk_fast21_tma_kernel, dim3(need < resident ? need : resident), dim3(32 * kLkWarpsPerCta), 0, st, p, n_streams,
                       max_features, p.work_counters, p.work_seq & 1, *p.tma_maps);
        else
            launch_pdl(lk_fast21_kernel, dim3(need < resident ? need : resident), dim3(32 * kLkWarpsPerCta), 0, st, p, n_streams, max_features,
                                                                                          p.work_counters, p.work_seq & 1);
    } else {
        const int smem = kLkWarpsPerCta * lk_warp_smem(p.win);
        ensure_dyn_smem(reinterpret_cast<const void*>(lk_kernel<0>), kLkWarpsPerCta * lk_warp_smem(kMaxWin));
        launch_pdl(lk_kernel<0>, grid, dim3(32 * kLkWarpsPerCta), smem, st, p);
    }
    if (launches) *launches += 1;
    return 0;
}

}  // namespace vs
