// Host-callable launchers of the vs_tracker CUDA kernels (one .cu per stage).
#pragma once
#include <cuda.h>   // CUtensorMap (type only; the encode function is fetched through the runtime)

#include "vs_common.cuh"

namespace vs {

void set_error(const char* fmt, ...);

// ---- clahe.cu (K1, K2) ----
void launch_clahe(const uint8_t* src, int src_pitch, int64_t src_stride, uint8_t* dst, int dst_pitch,
                  int64_t dst_stride, int W, int H, int n_streams, uint8_t* luts, cudaStream_t st,
                  int64_t* launches);
void launch_clahe_lut_only(const uint8_t* src, int src_pitch, int64_t src_stride, int W, int H, int n_streams,
                           uint8_t* luts, cudaStream_t st);
void launch_clahe_apply_only(const uint8_t* src, int src_pitch, int64_t src_stride, uint8_t* dst, int dst_pitch,
                             int64_t dst_stride, int W, int H, int n_streams, const uint8_t* luts,
                             cudaStream_t st);
void launch_copy_image(const uint8_t* src, int src_pitch, int64_t src_stride, uint8_t* dst, int dst_pitch,
                       int64_t dst_stride, int W, int H, int n_streams, cudaStream_t st, int64_t* launches);

// ---- pyramid.cu (K3) ----
void launch_pyr_down(const uint8_t* pyr_base, const PyrDesc& pd, int level_from, int n_streams, cudaStream_t st,
                     int64_t* launches);

// REFLECT_101 border (kPyrPadX / kPyrPadY) of levels 0 .. n_levels-1, two launches (columns, then rows)
void launch_pyr_pad(uint8_t* pyr_base, const PyrDesc& pd, int n_levels, int n_streams, cudaStream_t st, int64_t* launches);

// ---- lk.cu (K9) ----
struct LkParams {
    const uint8_t* prev_pyr;
    const uint8_t* next_pyr;
    PyrDesc pd;
    int n_levels;             // levels used = min(built prev, built next, maxLevel + 1)
    int win;
    int max_count;            // min(max(lk_iterations, 0), 100)
    double eps2;              // clamp(lk_eps, 0, 10)^2
    float min_eig_threshold;  // 1e-4
    const float2* pts;        // [n_streams][cap]
    const int* counts;        // [n_streams]
    int cap;
    float2* next_pts;         // [n_streams][cap]
    uint8_t* status;          // [n_streams][cap] raw LK status
    int* work_counters;       // [2] persistent-warp work queue heads (zeroed at create), alternated by work_seq
    int work_seq;             // launch sequence number of this handle
    // TMA staging (VS_LK_TMA=1, A/B only): tensor maps of both pyramid buffers and which of them is prev / next
    const struct LkTmaMaps* tma_maps;   // host pointer, may be null
    int prev_buf, next_buf;
};
// One 3-D tensor map (x bytes, rows, stream) per pyramid buffer, level and tile height (24 template rows / 32 search rows)
// over the level's PADDED block, so tile origins are never negative.
struct LkTmaMaps {
    CUtensorMap m[2][kMaxLevels][2];
};
// fills the maps of one handle (driver entry point cuTensorMapEncodeTiled through the runtime); 0 on success
int lk_build_tma_maps(LkTmaMaps* out, uint8_t* const d_pyr[2], const PyrDesc& pd, int n_streams);
int launch_lk(const LkParams& p, int n_streams, int max_features, cudaStream_t st, int64_t* launches);

// ---- shitomasi.cu (K5, K6) ----
struct StParams {
    const uint8_t* img;       // level 0 of the new pyramid
    int pitch;                // bytes, multiple of 16
    int64_t img_stride;       // bytes between two streams' images
    int W, H;
    int nx, ny, words_per_strip, rows_per_strip;   // strip grid (filled by launch_shi_tomasi)
    int skip_masked_rows;     // walk only the row steps that can reach an unmasked pixel (filled by launch_shi_tomasi)
    int img_padded;           // img is a pyramid level stored with its REFLECT_101 border (>= 8 columns each side): frame path
    const int* want;          // [n_streams] corners wanted; streams with want <= 0 are skipped (may be null)
    // analytic mask = base minus the min_dist discs of the kept features
    const int* kn;            // [n_streams] kept features (may be null = none)
    const float2* k_pts;      // [n_streams][cap]
    int cap;
    int min_dist;
    const int* disc_hw;       // [min_dist + 1] cv::circle row half widths
    int base_mode;            // 0 all pixels, 1 auto fisheye circle, 2 mask image
    int fcx, fcy, fr;         // fisheye circle (feature_tracker.cpp:50-58)
    const int* fish_hw;       // [fr + 1]
    const uint8_t* base_mask; // [H][mask_pitch] (base_mode 2)
    int mask_pitch;
    unsigned int* max_enc;    // [n_streams] order-preserving encoding of the masked max
    unsigned long long* cand; // [n_streams][cand_cap] (eig bits << 32 | y*W + x)
    int* cand_count;          // [n_streams]
    int cand_cap;
    // two-tier frame path (st_scan_kernel -> st_refine_kernel): records of the pixels that may be candidates / the
    // masked maximum, and the running lower bound of that maximum; null = one-pass exact kernel
    unsigned long long* raw;  // [n_streams][cand_cap] (neighbours' max upper bound bits << 32 | flags | y*W + x)
    int* raw_count;           // [n_streams]
    unsigned int* l_enc;      // [n_streams] order-preserving encoding of a lower bound of the masked max
    float* eig_out;           // optional dense eig map [n_streams][H][W] (stage tests), else null
    uint8_t* mask_out;        // optional mask image [n_streams][H][mask_out_pitch] (stage tests), else null
    int mask_out_pitch;
    int64_t mask_out_stride;
    double quality;           // 0.01
};
void launch_shi_tomasi(const StParams& p, int n_streams, cudaStream_t st, int64_t* launches);

__host__ __device__ inline unsigned int float_order_enc(float f) {
#ifdef __CUDA_ARCH__
    const unsigned int b = __float_as_uint(f);
#else
    union { float f; unsigned int u; } cv; cv.f = f; const unsigned int b = cv.u;
#endif
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ inline float float_order_dec(unsigned int e) {
    const unsigned int b = (e & 0x80000000u) ? (e & 0x7fffffffu) : ~e;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    union { float f; unsigned int u; } cv; cv.u = b; return cv.f;
#endif
}

// ---- glue.cu (K8, K10, K7, K11) ----
struct StreamState;   // defined in glue.cuh
struct GlueParams;

}  // namespace vs
