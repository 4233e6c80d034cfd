// K3 — Gaussian pyramid level (cv::pyrDown inside cv::buildOpticalFlowPyramid), bit-exact:
//   dst(x,y) = (sum_{i,j in [-2,2]} k_i k_j src(refl101(2x+i), refl101(2y+j)) + 128) >> 8, k=[1 4 6 4 1].
// Replaces the pyramid half of buildOpticalFlowPyramid at
// /root/reference/src/frontend/feature_tracker.cpp:123 (spec: SURVEY.md S2).  The Scharr derivative
// images the reference also stores (4 B/px) are never materialised: the LK kernel derives them
// from the staged u8 tile (lk.cu), which removes 4P bytes of writes and the matching reads per frame.
//
// Layout of the work: one warp walks down a strip of the destination level.  Per source row every lane
// loads 8 consecutive source pixels with one 64-bit load (256 coalesced bytes per warp), gets the two
// pixels to its left and the one to its right from its neighbour lanes by shuffle, and forms its four
// horizontal 5-tap sums with DP4A; the vertical 5-tap runs over a register window of the last five
// horizontal rows.  Each lane writes 4 destination pixels per row with one 32-bit store.  Every source
// byte is read from HBM/L2 once (the 3 halo rows per strip come from L2) and nothing is staged in
// shared memory.
#include "kernels.h"

namespace vs {

constexpr int kPdWarps = 4;        // warps per CTA, each owns one strip
constexpr int kPdLanes = 30;       // productive lanes per warp (lanes 0 and 31 only feed their neighbours)

__device__ __forceinline__ int dp4a_uu(unsigned a, unsigned b, int c) {
    int d;
    asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

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

    // lane l owns destination pixels 4*g .. 4*g+3 with g = ix*groups_per_strip + l - 1, i.e. source
    // pixels 8*g .. 8*g+7; lanes 0 / 31 (and those past the strip) only supply neighbours
    const int g = ix * groups_per_strip + lane - 1;
    const int ngroups = (dw + 3) >> 2;
    const bool productive = lane >= 1 && lane <= groups_per_strip && g < ngroups;
    const int sx0 = 8 * g;
    // a lane can use the aligned 64-bit load when its 8 pixels are inside the row; it can take its right
    // neighbour's pixels by shuffle when that lane's 8 pixels are inside too
    const bool inside = sx0 >= 0 && sx0 + 8 <= sw;
    const bool left_edge = (sx0 == 0);                                // s(-2), s(-1) by reflection
    const bool generic = productive && !(inside && sx0 + 16 <= sw);   // taps reach the right border
    const int y0 = iy * rows_per_strip;
    const int y1 = min(y0 + rows_per_strip, dh);

    // raw 8 source pixels of one source row (issued early so that several rows are in flight per warp)
    auto fetch = [&](int sy) -> uint2 {
        if (!inside) return make_uint2(0u, 0u);
        return __ldg(reinterpret_cast<const uint2*>(img + (int64_t)reflect101(sy, sh) * spitch + sx0));
    };
    // horizontal 5-tap sums of one source row for this lane's four destination pixels
    auto hrow = [&](int sy, uint2 raw, int (&h)[4]) {
        uint32_t lo = raw.x, hi = raw.y;
        uint32_t lw = __shfl_up_sync(full, hi, 1);       // source pixels sx0-4 .. sx0-1
        uint32_t rw = __shfl_down_sync(full, lo, 1);     // source pixels sx0+8 .. sx0+11
        if (generic) {
            // the lane(s) at the right border: REFLECT_101 byte loads for every tap
            const uint8_t* row = img + (int64_t)reflect101(sy, sh) * spitch;
            lo = hi = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                lo |= (uint32_t)__ldg(row + reflect101(sx0 + k, sw)) << (8 * k);
                hi |= (uint32_t)__ldg(row + reflect101(sx0 + 4 + k, sw)) << (8 * k);
            }
            rw = (uint32_t)__ldg(row + reflect101(sx0 + 8, sw));
            lw = (uint32_t)__ldg(row + reflect101(sx0 - 2, sw)) << 16 | (uint32_t)__ldg(row + reflect101(sx0 - 1, sw)) << 24;
        } else if (left_edge) {
            lw = __byte_perm(lo, 0, 0x1200);             // (.., .., s(2), s(1)) in the two high bytes
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
        uint32_t packed = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int v = a[k] + e[k] + 4 * (b[k] + d4[k]) + 6 * c[k];
            packed |= (uint32_t)((v + 128) >> 8) << (8 * k);
        }
        if (productive && y < y1) {
            uint8_t* d = out + (int64_t)y * dpitch + 4 * g;
            if (4 * g + 3 < dw) *reinterpret_cast<uint32_t*>(d) = packed;
            else
                for (int k = 0; k < 4 && 4 * g + k < dw; ++k) d[k] = (uint8_t)(packed >> (8 * k));
        }
    };

    int h0[4], h1[4], h2[4], h3[4], h4[4], h5[4], h6[4], h7[4], h8[4], h9[4], h10[4];
    {
        const uint2 r0 = fetch(2 * y0 - 2), r1 = fetch(2 * y0 - 1), r2 = fetch(2 * y0);
        hrow(2 * y0 - 2, r0, h0);
        hrow(2 * y0 - 1, r1, h1);
        hrow(2 * y0, r2, h2);
    }
    // four destination rows (eight source rows in flight) per trip
#pragma unroll 1
    for (int y = y0; y < y1; y += 4) {
        const int sy = 2 * y;
        const uint2 r3 = fetch(sy + 1), r4 = fetch(sy + 2), r5 = fetch(sy + 3), r6 = fetch(sy + 4);
        const uint2 r7 = fetch(sy + 5), r8 = fetch(sy + 6), r9 = fetch(sy + 7), r10 = fetch(sy + 8);
        hrow(sy + 1, r3, h3);
        hrow(sy + 2, r4, h4);
        emit(y, h0, h1, h2, h3, h4);
        hrow(sy + 3, r5, h5);
        hrow(sy + 4, r6, h6);
        emit(y + 1, h2, h3, h4, h5, h6);
        hrow(sy + 5, r7, h7);
        hrow(sy + 6, r8, h8);
        emit(y + 2, h4, h5, h6, h7, h8);
        hrow(sy + 7, r9, h9);
        hrow(sy + 8, r10, h10);
        emit(y + 3, h6, h7, h8, h9, h10);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            h0[k] = h8[k];
            h1[k] = h9[k];
            h2[k] = h10[k];
        }
    }
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

void launch_pyr_down(const uint8_t* pyr_base, const PyrDesc& pd, int level_from, int n_streams, cudaStream_t st,
                     int64_t* launches) {
    const LevelDesc& a = pd.lv[level_from];
    const LevelDesc& b = pd.lv[level_from + 1];
    const int ngroups = (b.w + 3) / 4;
    const int nx = (ngroups + kPdLanes - 1) / kPdLanes;
    const int gps = (ngroups + nx - 1) / nx;
    // short strips on the small levels (more warps, shorter dependent chains); each strip re-reads 3 source rows
    const int rows = b.h >= 192 ? 16 : 8;
    const int ny = (b.h + rows - 1) / rows;
    dim3 grid((nx * ny + kPdWarps - 1) / kPdWarps, n_streams);
    launch_pdl(pyr_down_kernel, grid, dim3(32 * kPdWarps), 0, st, pyr_base + a.offset, a.w, a.h, a.pitch,
                                                    const_cast<uint8_t*>(pyr_base) + b.offset, b.w, b.h, b.pitch,
                                                    pd.stream_bytes, nx, ny, gps, rows);
    if (launches) *launches += 1;
}

}  // namespace vs
