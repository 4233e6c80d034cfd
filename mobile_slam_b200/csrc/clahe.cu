// K1/K2 — CLAHE(clipLimit 3.0, 8x8 tiles), bit-exact to cv::CLAHE::apply on CV_8UC1.
// Replaces the clahe_->apply call at /root/reference/src/frontend/feature_tracker.cpp:102-106.
// Arithmetic spec: SURVEY.md S1 (integer histogram/clip/redistribute, float32 LUT scale and
// bilinear LUT blend with no FMA contraction, round-half-even).
#include "kernels.h"

namespace vs {

// ---------------------------------------------------------------------------------------------
// K1: one CTA per (tile, stream): 256-bin histogram -> clip -> redistribute -> scan -> LUT.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) clahe_lut_kernel(const uint8_t* __restrict__ src, int pitch,
                                                        int64_t img_stride, int W, int H, int tw, int th,
                                                        int clip, float lut_scale,
                                                        uint8_t* __restrict__ luts) {
    __shared__ unsigned int hist[8][256];
    __shared__ unsigned int warp_tot[8];
    __shared__ unsigned int s_clipped;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tile = blockIdx.x, s = blockIdx.y;
    const int tx = tile % kClaheTiles, ty = tile / kClaheTiles;
    const uint8_t* img = src + (int64_t)s * img_stride;
#pragma unroll
    for (int k = 0; k < 8; ++k) hist[k][tid] = 0;
    if (tid == 0) s_clipped = 0;
    __syncthreads();

    const int x0 = tx * tw, y0 = ty * th;
    const bool interior = (x0 + tw <= W) && (y0 + th <= H);
    if (interior && ((tw & 3) == 0) && ((pitch & 3) == 0)) {
        // fast path: 4 pixels per 32-bit load, tile rows are 4-byte aligned (x0 = tx*tw)
        const int wq = tw >> 2;
        for (int idx = tid; idx < wq * th; idx += 256) {
            const int y = idx / wq, xq = idx - y * wq;
            const uint32_t v = *reinterpret_cast<const uint32_t*>(img + (int64_t)(y0 + y) * pitch + x0 + 4 * xq);
            atomicAdd(&hist[warp][v & 255u], 1u);
            atomicAdd(&hist[warp][(v >> 8) & 255u], 1u);
            atomicAdd(&hist[warp][(v >> 16) & 255u], 1u);
            atomicAdd(&hist[warp][v >> 24], 1u);
        }
    } else {
        // general path: the image is virtually padded bottom/right with REFLECT_101 to 8x8 tiles
        for (int idx = tid; idx < tw * th; idx += 256) {
            const int y = idx / tw, x = idx - y * tw;
            const int gx = reflect101(x0 + x, W), gy = reflect101(y0 + y, H);
            atomicAdd(&hist[warp][img[(int64_t)gy * pitch + gx]], 1u);
        }
    }
    __syncthreads();
    int h = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) h += (int)hist[k][tid];

    // clip and count the excess
    int excess = h > clip ? h - clip : 0;
    h = h > clip ? clip : h;
    unsigned int e = (unsigned int)excess;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
    if (lane == 0 && e) atomicAdd(&s_clipped, e);
    __syncthreads();
    const int clipped = (int)s_clipped;
    h += clipped >> 8;                     // clipped / 256
    const int residual = clipped & 255;    // clipped % 256
    if (residual) {
        int step = 256 / residual;
        if (step < 1) step = 1;
        // serial loop "for (i = 0; i < 256 && residual > 0; i += step, residual--) hist[i]++"
        if ((tid % step) == 0 && (tid / step) < residual) h += 1;
    }
    // inclusive scan over the 256 bins
    unsigned int v = (unsigned int)h;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned int n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    if (lane == 31) warp_tot[warp] = v;
    __syncthreads();
    unsigned int base = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k)
        if (k < warp) base += warp_tot[k];
    const float sum = (float)(int)(v + base);
    int q = __float2int_rn(__fmul_rn(sum, lut_scale));
    q = q < 0 ? 0 : (q > 255 ? 255 : q);
    luts[((int64_t)s * (kClaheTiles * kClaheTiles) + tile) * 256 + tid] = (uint8_t)q;
}

// ---------------------------------------------------------------------------------------------
// K2: bilinear blend of the four neighbouring tile LUTs.  One CTA = a band of kBandRows rows.
// ---------------------------------------------------------------------------------------------
constexpr int kBandRows = 16;

__device__ __forceinline__ uint32_t clahe_px(const uint8_t* __restrict__ lut_lo,
                                             const uint8_t* __restrict__ lut_hi, int tx1, int tx2,
                                             float xa, float xa1, float ya, float ya1, uint32_t v) {
    const float p1 = (float)lut_lo[tx1 * 256 + v];
    const float p2 = (float)lut_lo[tx2 * 256 + v];
    const float q1 = (float)lut_hi[tx1 * 256 + v];
    const float q2 = (float)lut_hi[tx2 * 256 + v];
    const float top = __fadd_rn(__fmul_rn(p1, xa1), __fmul_rn(p2, xa));
    const float bot = __fadd_rn(__fmul_rn(q1, xa1), __fmul_rn(q2, xa));
    const float r = __fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya));
    int q = __float2int_rn(r);
    q = q < 0 ? 0 : (q > 255 ? 255 : q);
    return (uint32_t)q;
}

__global__ void __launch_bounds__(256) clahe_apply_kernel(const uint8_t* __restrict__ src, int src_pitch,
                                                          int64_t src_stride, uint8_t* __restrict__ dst,
                                                          int dst_pitch, int64_t dst_stride, int W, int H,
                                                          float inv_tw, float inv_th,
                                                          const uint8_t* __restrict__ luts) {
    __shared__ __align__(16) uint8_t s_lut[kClaheTiles * kClaheTiles * 256];
    const int tid = threadIdx.x;
    const int s = blockIdx.y;
    const int y_begin = blockIdx.x * kBandRows;
    const int y_end = min(y_begin + kBandRows, H);
    // tile rows touched by this band
    const int ty_first = max((int)floorf(__fsub_rn(__fmul_rn((float)y_begin, inv_th), 0.5f)), 0);
    const int ty_last = min((int)floorf(__fsub_rn(__fmul_rn((float)(y_end - 1), inv_th), 0.5f)) + 1, kClaheTiles - 1);
    {
        const uint4* g = reinterpret_cast<const uint4*>(luts + (int64_t)s * (kClaheTiles * kClaheTiles * 256));
        uint4* d = reinterpret_cast<uint4*>(s_lut);
        const int lo = ty_first * (kClaheTiles * 256 / 16), hi = (ty_last + 1) * (kClaheTiles * 256 / 16);
        for (int i = lo + tid; i < hi; i += 256) d[i] = g[i];
    }
    __syncthreads();
    const uint8_t* img = src + (int64_t)s * src_stride;
    uint8_t* out = dst + (int64_t)s * dst_stride;
    const int wq = (W + 3) >> 2;
    const int rows = y_end - y_begin;
    for (int idx = tid; idx < wq * rows; idx += 256) {
        const int r = idx / wq, xq = idx - r * wq;
        const int y = y_begin + r, x = xq << 2;
        const float tyf = __fsub_rn(__fmul_rn((float)y, inv_th), 0.5f);
        const int ty1r = (int)floorf(tyf);
        const float ya = __fsub_rn(tyf, (float)ty1r), ya1 = __fsub_rn(1.0f, ya);
        const int ty1 = max(ty1r, 0), ty2 = min(ty1r + 1, kClaheTiles - 1);
        const uint8_t* lut_lo = s_lut + ty1 * (kClaheTiles * 256);
        const uint8_t* lut_hi = s_lut + ty2 * (kClaheTiles * 256);
        const uint8_t* sp = img + (int64_t)y * src_pitch + x;
        uint8_t* dp = out + (int64_t)y * dst_pitch + x;
        uint32_t v4;
        const bool full = (x + 3 < W);
        if (full) v4 = *reinterpret_cast<const uint32_t*>(sp);
        else {
            v4 = 0;
            for (int k = 0; k < 4 && x + k < W; ++k) v4 |= (uint32_t)sp[k] << (8 * k);
        }
        uint32_t o4 = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float txf = __fsub_rn(__fmul_rn((float)(x + k), inv_tw), 0.5f);
            const int tx1r = (int)floorf(txf);
            const float xa = __fsub_rn(txf, (float)tx1r), xa1 = __fsub_rn(1.0f, xa);
            const int tx1 = max(tx1r, 0), tx2 = min(tx1r + 1, kClaheTiles - 1);
            o4 |= clahe_px(lut_lo, lut_hi, tx1, tx2, xa, xa1, ya, ya1, (v4 >> (8 * k)) & 255u) << (8 * k);
        }
        if (full) *reinterpret_cast<uint32_t*>(dp) = o4;
        else
            for (int k = 0; k < 4 && x + k < W; ++k) dp[k] = (uint8_t)(o4 >> (8 * k));
    }
}

// plain strided copy used when equalize == 0 (level 0 = the raw frame)
__global__ void __launch_bounds__(256) copy_image_kernel(const uint8_t* __restrict__ src, int src_pitch,
                                                         int64_t src_stride, uint8_t* __restrict__ dst,
                                                         int dst_pitch, int64_t dst_stride, int W, int H) {
    const int s = blockIdx.y;
    const int wq = (W + 3) >> 2;
    for (int idx = blockIdx.x * 256 + threadIdx.x; idx < wq * H; idx += gridDim.x * 256) {
        const int y = idx / wq, x = (idx - y * wq) << 2;
        const uint8_t* sp = src + (int64_t)s * src_stride + (int64_t)y * src_pitch + x;
        uint8_t* dp = dst + (int64_t)s * dst_stride + (int64_t)y * dst_pitch + x;
        if (x + 3 < W && ((src_pitch & 3) == 0) && ((reinterpret_cast<uintptr_t>(sp) & 3) == 0))
            *reinterpret_cast<uint32_t*>(dp) = *reinterpret_cast<const uint32_t*>(sp);
        else
            for (int k = 0; k < 4 && x + k < W; ++k) dp[k] = sp[k];
    }
}

void launch_clahe(const uint8_t* src, int src_pitch, int64_t src_stride, uint8_t* dst, int dst_pitch,
                  int64_t dst_stride, int W, int H, int n_streams, uint8_t* luts, cudaStream_t st,
                  int64_t* launches) {
    const int We = (W + kClaheTiles - 1) / kClaheTiles * kClaheTiles;
    const int He = (H + kClaheTiles - 1) / kClaheTiles * kClaheTiles;
    const int tw = We / kClaheTiles, th = He / kClaheTiles;
    const int area = tw * th;
    int clip = (int)(3.0 * area / 256);   // clipLimit * tileSizeTotal / histSize
    if (clip < 1) clip = 1;
    const float lut_scale = 255.0f / (float)area;
    clahe_lut_kernel<<<dim3(kClaheTiles * kClaheTiles, n_streams), 256, 0, st>>>(
        src, src_pitch, src_stride, W, H, tw, th, clip, lut_scale, luts);
    clahe_apply_kernel<<<dim3((H + kBandRows - 1) / kBandRows, n_streams), 256, 0, st>>>(
        src, src_pitch, src_stride, dst, dst_pitch, dst_stride, W, H, 1.0f / (float)tw, 1.0f / (float)th, luts);
    if (launches) *launches += 2;
}

void launch_clahe_lut_only(const uint8_t* src, int src_pitch, int64_t src_stride, int W, int H, int n_streams,
                           uint8_t* luts, cudaStream_t st) {
    const int We = (W + kClaheTiles - 1) / kClaheTiles * kClaheTiles;
    const int He = (H + kClaheTiles - 1) / kClaheTiles * kClaheTiles;
    const int tw = We / kClaheTiles, th = He / kClaheTiles;
    int clip = (int)(3.0 * (tw * th) / 256);
    if (clip < 1) clip = 1;
    clahe_lut_kernel<<<dim3(kClaheTiles * kClaheTiles, n_streams), 256, 0, st>>>(
        src, src_pitch, src_stride, W, H, tw, th, clip, 255.0f / (float)(tw * th), luts);
}

void launch_clahe_apply_only(const uint8_t* src, int src_pitch, int64_t src_stride, uint8_t* dst, int dst_pitch,
                             int64_t dst_stride, int W, int H, int n_streams, const uint8_t* luts,
                             cudaStream_t st) {
    const int We = (W + kClaheTiles - 1) / kClaheTiles * kClaheTiles;
    const int He = (H + kClaheTiles - 1) / kClaheTiles * kClaheTiles;
    const int tw = We / kClaheTiles, th = He / kClaheTiles;
    clahe_apply_kernel<<<dim3((H + kBandRows - 1) / kBandRows, n_streams), 256, 0, st>>>(
        src, src_pitch, src_stride, dst, dst_pitch, dst_stride, W, H, 1.0f / (float)tw, 1.0f / (float)th, luts);
}

void launch_copy_image(const uint8_t* src, int src_pitch, int64_t src_stride, uint8_t* dst, int dst_pitch,
                       int64_t dst_stride, int W, int H, int n_streams, cudaStream_t st, int64_t* launches) {
    const int wq = (W + 3) >> 2;
    int blocks = (wq * H + 255) / 256;
    if (blocks > 1024) blocks = 1024;
    copy_image_kernel<<<dim3(blocks, n_streams), 256, 0, st>>>(src, src_pitch, src_stride, dst, dst_pitch,
                                                               dst_stride, W, H);
    if (launches) *launches += 1;
}

}  // namespace vs
