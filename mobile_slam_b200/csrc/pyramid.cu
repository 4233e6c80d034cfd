// K3 — Gaussian pyramid level (cv::pyrDown inside cv::buildOpticalFlowPyramid), bit-exact:
//   dst(x,y) = (sum_{i,j in [-2,2]} k_i k_j src(refl101(2x+i), refl101(2y+j)) + 128) >> 8, k=[1 4 6 4 1].
// Replaces the pyramid half of buildOpticalFlowPyramid at
// /root/reference/src/frontend/feature_tracker.cpp:123 (spec: SURVEY.md S2).  The Scharr derivative
// images the reference also stores (4 B/px) are never materialised: the LK kernel derives them
// from the staged u8 tile (lk.cu), which removes 4P bytes of writes and the matching reads per frame.
//
// Layout of the work: one warp walks down a strip of the destination level.  Per source row every lane
// loads 8 consecutive source pixels with one 64-bit load (256 coalesced bytes per warp), gets the four
// pixels to its left and the four to its right from its neighbour lanes by shuffle (lane 0 / lane 31 load
// those words themselves, so all 32 lanes produce output), and forms its four horizontal 5-tap sums with
// DP4A; the vertical 5-tap runs over a register window of the last five horizontal rows.  Each lane writes
// 4 destination pixels per row with one 32-bit store.  Every source byte is read from HBM/L2 once (the 3
// halo rows per strip come from L2) and nothing is staged in shared memory.
//
// Borders.  Levels live in padded blocks (vs_common.cuh), so loads a few bytes outside a row are legal —
// but the source level's border is not written yet when this kernel runs, so the values must not come
// from there: the two pixels left of column 0 are a byte permutation of the first word, and the two right
// of column sw-1 are patched into the registers of the ONE lane that holds the end of the row (its position
// inside the lane's 12-byte window is sw mod 8, a template parameter, so the patch is one or two PRMTs).
// No lane ever takes a byte-wise path, and the kernel is a few hundred instructions instead of two thousand.
#include "kernels.h"

namespace vs {

static_assert(kPyrFirstPaddedLevel == 0 && kPyrPadX >= 16 && kPyrPadY >= 10, "pyr_down reads up to 15 bytes / 8 rows outside a level");

constexpr int kPdWarps = 4;        // warps per CTA, each owns one strip
constexpr int kPdLanes = 32;       // destination groups (4 pixels) per strip row

__device__ __forceinline__ int dp4a_uu(unsigned a, unsigned b, int c) {
    int d;
    asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// R = sw - 8 * ((sw - 1) / 8) in 1..8: the lane holding the last source pixel has R valid pixels in (lo, hi); window
// bytes R and R + 1 of (lo, hi, rw) must read s(sw-2) and s(sw-3), which are window bytes R - 2 and R - 3 (lw = bytes -4..-1)
template <int R>
__device__ __forceinline__ void patch_right_border(uint32_t lw, uint32_t& lo, uint32_t& hi, uint32_t& rw) {
    if constexpr (R == 1) lo = __byte_perm(lo, lw, 0x3670);                       // lo[1] = lw[3], lo[2] = lw[2]
    else if constexpr (R == 2) lo = __byte_perm(lo, lw, 0x7010);                  // lo[2] = lo[0], lo[3] = lw[3]
    else if constexpr (R == 3) { hi = __byte_perm(hi, lo, 0x3214); lo = __byte_perm(lo, 0, 0x1210); }   // hi[0] = lo[0], lo[3] = lo[1]
    else if constexpr (R == 4) hi = __byte_perm(hi, lo, 0x3256);                  // hi[0] = lo[2], hi[1] = lo[1]
    else if constexpr (R == 5) hi = __byte_perm(hi, lo, 0x3670);                  // hi[1] = lo[3], hi[2] = lo[2]
    else if constexpr (R == 6) hi = __byte_perm(hi, lo, 0x7010);                  // hi[2] = hi[0], hi[3] = lo[3]
    else if constexpr (R == 7) { rw = __byte_perm(rw, hi, 0x3214); hi = __byte_perm(hi, 0, 0x1210); }   // rw[0] = hi[0], hi[3] = hi[1]
    else rw = __byte_perm(rw, hi, 0x3256);                                        // rw[0] = hi[2], rw[1] = hi[1]
}

template <int R>
__global__ void __launch_bounds__(32 * kPdWarps) pyr_down_kernel(const uint8_t* __restrict__ src, int sw, int sh,
                                                                 int spitch, uint8_t* __restrict__ dst, int dw, int dh,
                                                                 int dpitch, int64_t stream_stride, int nx, int ny,
                                                                 int groups_per_strip, int rows_per_strip) {
    pdl_entry();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int strip = blockIdx.x * kPdWarps + warp;
    if (strip >= nx * ny) return;
    const int ix = strip % nx, iy = strip / nx;
    const int s = blockIdx.y;
    const uint8_t* img = src + (int64_t)s * stream_stride;
    uint8_t* out = dst + (int64_t)s * stream_stride;
    const unsigned full = 0xffffffffu;

    // lane l owns destination pixels 4g .. 4g+3 with g = ix * groups_per_strip + l, i.e. source pixels 8g .. 8g+7
    const int g = ix * groups_per_strip + lane;
    const int ngroups = (dw + 3) >> 2;
    const bool productive = lane < groups_per_strip && g < ngroups;
    const int sx0 = 8 * g;
    const bool live = sx0 < sw + 8;                     // lanes further right neither produce nor feed a neighbour
    const bool first = lane == 0, last = lane == 31;
    const bool left_edge = sx0 == 0;
    const int g_end = (sw - 1) >> 3;                    // group holding the last source pixel
    const bool border_warp = g_end >= ix * groups_per_strip && g_end < ix * groups_per_strip + 32;     // warp-uniform
    const bool border_lane = g == g_end;
    const int ext_dx = first ? -4 : 8;                  // lane 0: the word left of its group, lane 31: the word right of it
    const int y0 = iy * rows_per_strip;
    const int y1 = min(y0 + rows_per_strip, dh);

    struct Raw { uint2 v; uint32_t ext; };
    // the 8 source pixels of one source row (+ the strip's outer words), issued early so several rows are in flight
    auto fetch = [&](int sy) -> Raw {
        Raw r;
        r.v = make_uint2(0u, 0u);
        r.ext = 0u;
        const uint8_t* row = img + (int64_t)reflect101(sy, sh) * spitch + sx0;
        if (live) r.v = __ldg(reinterpret_cast<const uint2*>(row));
        if (live && (first || last)) r.ext = __ldg(reinterpret_cast<const uint32_t*>(row + ext_dx));
        return r;
    };
    // horizontal 5-tap sums of one source row for this lane's four destination pixels
    auto hrow = [&](const Raw& raw, int (&h)[4]) {
        uint32_t lo = raw.v.x, hi = raw.v.y;
        uint32_t lw = __shfl_up_sync(full, hi, 1);       // source pixels sx0-4 .. sx0-1
        uint32_t rw = __shfl_down_sync(full, lo, 1);     // source pixels sx0+8 .. sx0+11
        if (first) lw = raw.ext;
        if (last) rw = raw.ext;
        if (left_edge) lw = __byte_perm(lo, 0, 0x1200);  // (.., .., s(2), s(1)) in the two high bytes
        if (border_warp) {
            uint32_t plo = lo, phi = hi, prw = rw;
            patch_right_border<R>(lw, plo, phi, prw);
            if (border_lane) { lo = plo; hi = phi; rw = prw; }
        }
        // windows of 4 source pixels starting at sx0-2, sx0, sx0+2, sx0+4 (taps 1 4 6 4) and the 5th tap
        const uint32_t wm2 = __funnelshift_r(lw, lo, 16), w2 = __funnelshift_r(lo, hi, 16), w6 = __funnelshift_r(hi, rw, 16);
        const unsigned C = 0x04060401u, ONE = 0x00000001u;
        h[0] = dp4a_uu(w2, ONE, dp4a_uu(wm2, C, 0));
        h[1] = dp4a_uu(hi, ONE, dp4a_uu(lo, C, 0));
        h[2] = dp4a_uu(w6, ONE, dp4a_uu(w2, C, 0));
        h[3] = dp4a_uu(rw, ONE, dp4a_uu(hi, C, 0));
    };
    auto emit = [&](int y, const int (&a)[4], const int (&b)[4], const int (&c)[4], const int (&d4)[4], const int (&e)[4]) {
        uint32_t v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = (uint32_t)(a[k] + e[k] + 128 + 4 * (b[k] + d4[k]) + 6 * c[k]);     // < 65536
        // byte 1 of each sum = (sum + 128) >> 8
        const uint32_t packed = __byte_perm(__byte_perm(v[0], v[1], 0x0051), __byte_perm(v[2], v[3], 0x0051), 0x5410);
        if (productive && y < y1) {
            uint8_t* d = out + (int64_t)y * dpitch + 4 * g;
            if (4 * g + 3 < dw) *reinterpret_cast<uint32_t*>(d) = packed;
            else
                for (int k = 0; k < 4 && 4 * g + k < dw; ++k) d[k] = (uint8_t)(packed >> (8 * k));
        }
    };

    int h0[4], h1[4], h2[4], h3[4], h4[4], h5[4], h6[4], h7[4], h8[4], h9[4], h10[4];
    {
        const Raw r0 = fetch(2 * y0 - 2), r1 = fetch(2 * y0 - 1), r2 = fetch(2 * y0);
        hrow(r0, h0);
        hrow(r1, h1);
        hrow(r2, h2);
    }
    // four destination rows (eight source rows in flight) per trip
#pragma unroll 1
    for (int y = y0; y < y1; y += 4) {
        const int sy = 2 * y;
        const Raw r3 = fetch(sy + 1), r4 = fetch(sy + 2), r5 = fetch(sy + 3), r6 = fetch(sy + 4);
        const Raw r7 = fetch(sy + 5), r8 = fetch(sy + 6), r9 = fetch(sy + 7), r10 = fetch(sy + 8);
        hrow(r3, h3);
        hrow(r4, h4);
        emit(y, h0, h1, h2, h3, h4);
        hrow(r5, h5);
        hrow(r6, h6);
        emit(y + 1, h2, h3, h4, h5, h6);
        hrow(r7, h7);
        hrow(r8, h8);
        emit(y + 2, h4, h5, h6, h7, h8);
        hrow(r9, h9);
        hrow(r10, h10);
        emit(y + 3, h6, h7, h8, h9, h10);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            h0[k] = h8[k];
            h1[k] = h9[k];
            h2[k] = h10[k];
        }
    }
}

// levels narrower than 16 pixels (only reachable with tiny images / LK windows): one thread per destination pixel,
// the defining formula with REFLECT_101 on every tap
__global__ void __launch_bounds__(128) pyr_down_small_kernel(const uint8_t* __restrict__ src, int sw, int sh, int spitch,
                                                             uint8_t* __restrict__ dst, int dw, int dh, int dpitch,
                                                             int64_t stream_stride) {
    pdl_entry();
    const int idx = blockIdx.x * 128 + threadIdx.x;
    if (idx >= dw * dh) return;
    const int y = idx / dw, x = idx - y * dw;
    const uint8_t* img = src + (int64_t)blockIdx.y * stream_stride;
    const int k5[5] = {1, 4, 6, 4, 1};
    int acc = 128;
#pragma unroll
    for (int j = -2; j <= 2; ++j) {
        const uint8_t* row = img + (int64_t)reflect101(2 * y + j, sh) * spitch;
        int hs = 0;
#pragma unroll
        for (int i = -2; i <= 2; ++i) hs += k5[i + 2] * (int)row[reflect101(2 * x + i, sw)];
        acc += k5[j + 2] * hs;
    }
    dst[(int64_t)blockIdx.y * stream_stride + (int64_t)y * dpitch + x] = (uint8_t)(acc >> 8);
}

// ---- REFLECT_101 border of levels kPyrFirstPaddedLevel and up (cv::buildOpticalFlowPyramid stores its levels padded the same way) -----
// One warp per row.  Columns first: the 48 bytes left of the row and everything from the last aligned word of the row to
// the end of the pitch, one aligned word per lane: the four reflected source bytes t-3 .. t sit in two aligned words of the
// row, which are funnel-shifted together and byte-reversed.  Then rows: pad row y is a 16-byte copy of the whole padded
// row reflect101(y), corners included.
// Measured alternatives (profiles/r2_experiments.json): one fused launch with 4 rows per warp (0.119 ms per C2 step
// against 0.102 for these two launches) and padding inside the producers' CTAs (pyr_down with one CTA per row band +
// clahe_apply, after a CTA barrier: 0.121 ms and +16 us single-stream latency).  The cost is that of ~600 K small,
// scattered row-end writes per step, not of the re-read of rows that have left L2.
constexpr int kPadWarps = 8;

__device__ __forceinline__ bool pad_locate(const PyrDesc& pd, int n_levels, int rows_of_level_extra, int idx, int& level, int& row) {
    for (int l = kPyrFirstPaddedLevel; l < n_levels; ++l) {
        const int n = rows_of_level_extra > 0 ? rows_of_level_extra : pd.lv[l].h;
        if (idx < n) { level = l; row = idx; return true; }
        idx -= n;
    }
    return false;
}

__global__ void __launch_bounds__(32 * kPadWarps) pyr_pad_x_kernel(uint8_t* __restrict__ pyr, PyrDesc pd, int n_levels) {
    pdl_entry();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int l, y;
    if (!pad_locate(pd, n_levels, 0, blockIdx.x * kPadWarps + warp, l, y)) return;
    const LevelDesc L = pd.lv[l];
    uint8_t* row = pyr + (int64_t)blockIdx.y * pd.stream_bytes + L.offset + (int64_t)y * L.pitch;
    constexpr int NL = kPyrPadX / 4;
    const int xr0 = L.w & ~3, nright = (L.pitch - kPyrPadX - xr0) >> 2;
    int x0;
    if (lane < NL) x0 = -kPyrPadX + 4 * lane;
    else if (lane - NL < nright) x0 = xr0 + 4 * (lane - NL);
    else return;
    const int t = x0 < 0 ? -x0 : 2 * L.w - 2 - x0;           // source column of the word's FIRST byte; then t-1, t-2, t-3
    uint32_t v;
    if (L.w >= 64 && (x0 < 0 || x0 >= L.w)) {
        const int a = (t - 3) & ~3;
        const uint32_t A = __ldg(reinterpret_cast<const uint32_t*>(row + a)), B = __ldg(reinterpret_cast<const uint32_t*>(row + a + 4));
        v = __byte_perm(__funnelshift_r(A, B, ((t - 3) & 3) * 8), 0, 0x0123);
    } else {                                                  // the word straddling the right edge, and tiny levels
        v = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) v |= (uint32_t)row[reflect101(x0 + b, L.w)] << (8 * b);     // in-range bytes map to themselves
    }
    *reinterpret_cast<uint32_t*>(row + x0) = v;
}

__global__ void __launch_bounds__(32 * kPadWarps) pyr_pad_y_kernel(uint8_t* __restrict__ pyr, PyrDesc pd, int n_levels) {
    pdl_entry();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int l, j;
    if (!pad_locate(pd, n_levels, 2 * kPyrPadY, blockIdx.x * kPadWarps + warp, l, j)) return;
    const LevelDesc L = pd.lv[l];
    const int y = j < kPyrPadY ? j - kPyrPadY : L.h + (j - kPyrPadY);
    uint8_t* base = pyr + (int64_t)blockIdx.y * pd.stream_bytes + L.offset - kPyrPadX;        // column -kPyrPadX of row 0
    const uint4* src = reinterpret_cast<const uint4*>(base + (int64_t)reflect101(y, L.h) * L.pitch);
    uint4* dst = reinterpret_cast<uint4*>(base + (int64_t)y * L.pitch);
    for (int c = lane; c < (L.pitch >> 4); c += 32) dst[c] = src[c];
}

void launch_pyr_pad(uint8_t* pyr_base, const PyrDesc& pd, int n_levels, int n_streams, cudaStream_t st, int64_t* launches) {
    int rows = 0;
    for (int l = kPyrFirstPaddedLevel; l < n_levels; ++l) rows += pd.lv[l].h;
    if (rows == 0) return;
    launch_pdl(pyr_pad_x_kernel, dim3((rows + kPadWarps - 1) / kPadWarps, n_streams), dim3(32 * kPadWarps), 0, st, pyr_base, pd, n_levels);
    launch_pdl(pyr_pad_y_kernel, dim3((2 * kPyrPadY * (n_levels - kPyrFirstPaddedLevel) + kPadWarps - 1) / kPadWarps, n_streams), dim3(32 * kPadWarps), 0, st,
               pyr_base, pd, n_levels);
    if (launches) *launches += 2;
}

template <int R>
static void launch_pyr_down_r(const uint8_t* pyr_base, const PyrDesc& pd, const LevelDesc& a, const LevelDesc& b, int n_streams,
                              cudaStream_t st) {
    const int ngroups = (b.w + 3) / 4;
    const int nx = (ngroups + kPdLanes - 1) / kPdLanes;
    const int gps = (ngroups + nx - 1) / nx;
    // short strips on the small levels (more warps, shorter dependent chains); each strip re-reads 3 source rows
    const int rows = b.h >= 192 ? 16 : 8;
    const int ny = (b.h + rows - 1) / rows;
    dim3 grid((nx * ny + kPdWarps - 1) / kPdWarps, n_streams);
    launch_pdl(pyr_down_kernel<R>, grid, dim3(32 * kPdWarps), 0, st, pyr_base + a.offset, a.w, a.h, a.pitch,
                                                       const_cast<uint8_t*>(pyr_base) + b.offset, b.w, b.h, b.pitch,
                                                       pd.stream_bytes, nx, ny, gps, rows);
}

void launch_pyr_down(const uint8_t* pyr_base, const PyrDesc& pd, int level_from, int n_streams, cudaStream_t st,
                     int64_t* launches) {
    const LevelDesc& a = pd.lv[level_from];
    const LevelDesc& b = pd.lv[level_from + 1];
    if (launches) *launches += 1;
    if (a.w < 16) {
        launch_pdl(pyr_down_small_kernel, dim3((b.w * b.h + 127) / 128, n_streams), dim3(128), 0, st, pyr_base + a.offset, a.w, a.h,
                   a.pitch, const_cast<uint8_t*>(pyr_base) + b.offset, b.w, b.h, b.pitch, pd.stream_bytes);
        return;
    }
    switch (a.w - 8 * ((a.w - 1) / 8)) {
        case 1: launch_pyr_down_r<1>(pyr_base, pd, a, b, n_streams, st); break;
        case 2: launch_pyr_down_r<2>(pyr_base, pd, a, b, n_streams, st); break;
        case 3: launch_pyr_down_r<3>(pyr_base, pd, a, b, n_streams, st); break;
        case 4: launch_pyr_down_r<4>(pyr_base, pd, a, b, n_streams, st); break;
        case 5: launch_pyr_down_r<5>(pyr_base, pd, a, b, n_streams, st); break;
        case 6: launch_pyr_down_r<6>(pyr_base, pd, a, b, n_streams, st); break;
        case 7: launch_pyr_down_r<7>(pyr_base, pd, a, b, n_streams, st); break;
        default: launch_pyr_down_r<8>(pyr_base, pd, a, b, n_streams, st); break;
    }
}

}  // namespace vs
