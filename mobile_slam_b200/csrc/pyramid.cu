// K3 — Gaussian pyramid level (cv::pyrDown inside cv::buildOpticalFlowPyramid), bit-exact:
//   dst(x,y) = (sum_{i,j in [-2,2]} k_i k_j src(refl101(2x+i), refl101(2y+j)) + 128) >> 8, k=[1 4 6 4 1].
// Replaces the pyramid half of buildOpticalFlowPyramid at
// /root/reference/src/frontend/feature_tracker.cpp:123 (spec: SURVEY.md S2).  The Scharr derivative
// images the reference also stores (4 B/px) are never materialised: the LK kernel derives them
// from the staged u8 tile (lk.cu), which removes 4P bytes of writes and the matching reads per frame.
#include "kernels.h"

namespace vs {

constexpr int kTileW = 64;   // output tile width
constexpr int kTileH = 16;   // output tile height
constexpr int kSrcW = 2 * kTileW + 3;   // 131
constexpr int kSrcH = 2 * kTileH + 3;   // 35
constexpr int kSrcPitch = 144;          // bytes, multiple of 16

__global__ void __launch_bounds__(256) pyr_down_kernel(const uint8_t* __restrict__ src, int sw, int sh,
                                                       int spitch, uint8_t* __restrict__ dst, int dw, int dh,
                                                       int dpitch, int64_t stream_stride) {
    __shared__ __align__(16) uint8_t s_src[kSrcH * kSrcPitch];
    __shared__ __align__(16) uint16_t s_h[kSrcH * kTileW];
    const int tid = threadIdx.x;
    const int s = blockIdx.z;
    const uint8_t* img = src + (int64_t)s * stream_stride;
    uint8_t* out = dst + (int64_t)s * stream_stride;
    const int ox0 = blockIdx.x * kTileW, oy0 = blockIdx.y * kTileH;
    const int sx0 = 2 * ox0 - 2, sy0 = 2 * oy0 - 2;

    const bool interior = (sx0 >= 0) && (sy0 >= 0) && (sx0 + kSrcPitch <= spitch) && (sy0 + kSrcH <= sh) &&
                          (sx0 + kSrcW <= sw);
    if (interior) {
        // sx0 = 128*bx - 2: read from the 16-byte aligned address 2 bytes... keep it simple: 4-byte
        // words starting at sx0 - 2 (aligned because sx0 = 2 (mod 4)) shifted into place.
        const int wq = (kSrcW + 2 + 3) / 4;   // words covering [sx0-2, sx0+kSrcW)
        for (int idx = tid; idx < kSrcH * wq; idx += 256) {
            const int r = idx / wq, q = idx - r * wq;
            const uint8_t* p = img + (int64_t)(sy0 + r) * spitch + (sx0 - 2) + 4 * q;
            const uint32_t v = *reinterpret_cast<const uint32_t*>(p);
            // destination column of byte 0 of this word is 4q-2
            const int c = 4 * q - 2;
            uint8_t* d = s_src + r * kSrcPitch;
            if (c >= 0) d[c] = (uint8_t)v;
            if (c + 1 >= 0) d[c + 1] = (uint8_t)(v >> 8);
            d[c + 2] = (uint8_t)(v >> 16);
            d[c + 3] = (uint8_t)(v >> 24);
        }
    } else {
        for (int idx = tid; idx < kSrcH * kSrcW; idx += 256) {
            const int r = idx / kSrcW, c = idx - r * kSrcW;
            const int gy = reflect101(sy0 + r, sh), gx = reflect101(sx0 + c, sw);
            s_src[r * kSrcPitch + c] = img[(int64_t)gy * spitch + gx];
        }
    }
    __syncthreads();
    // horizontal 5-tap at even columns: h(r, x) = s(2x) + 4 s(2x+1) + 6 s(2x+2) + 4 s(2x+3) + s(2x+4)
    for (int idx = tid; idx < kSrcH * kTileW; idx += 256) {
        const int r = idx / kTileW, x = idx - r * kTileW;
        const uint8_t* p = s_src + r * kSrcPitch + 2 * x;
        s_h[idx] = (uint16_t)(p[0] + 4 * (p[1] + p[3]) + 6 * p[2] + p[4]);
    }
    __syncthreads();
    // vertical 5-tap, 4 outputs per thread -> one 32-bit store
    {
        const int y = tid >> 4, x = (tid & 15) << 2;
        const int oy = oy0 + y, ox = ox0 + x;
        if (oy < dh && ox < dw) {
            uint32_t packed = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint16_t* p = s_h + (2 * y) * kTileW + x + k;
                const int v = p[0] + 4 * (p[kTileW] + p[3 * kTileW]) + 6 * p[2 * kTileW] + p[4 * kTileW];
                packed |= (uint32_t)((v + 128) >> 8) << (8 * k);
            }
            uint8_t* d = out + (int64_t)oy * dpitch + ox;
            if (ox + 3 < dw) *reinterpret_cast<uint32_t*>(d) = packed;
            else
                for (int k = 0; k < 4 && ox + k < dw; ++k) d[k] = (uint8_t)(packed >> (8 * k));
        }
    }
}

void launch_pyr_down(const uint8_t* pyr_base, const PyrDesc& pd, int level_from, int n_streams, cudaStream_t st,
                     int64_t* launches) {
    const LevelDesc& a = pd.lv[level_from];
    const LevelDesc& b = pd.lv[level_from + 1];
    dim3 grid((b.w + kTileW - 1) / kTileW, (b.h + kTileH - 1) / kTileH, n_streams);
    pyr_down_kernel<<<grid, 256, 0, st>>>(pyr_base + a.offset, a.w, a.h, a.pitch,
                                          const_cast<uint8_t*>(pyr_base) + b.offset, b.w, b.h, b.pitch,
                                          pd.stream_bytes);
    if (launches) *launches += 1;
}

}  // namespace vs
