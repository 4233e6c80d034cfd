// Shared device/host definitions for the vs_tracker kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vs {

constexpr int kMaxLevels = 6;      // maxLevel <= 5  (VIOEngine::setTrackingParams, vio_engine.cpp:369)
constexpr int kClaheTiles = 8;     // createCLAHE(3.0, Size(8,8)), feature_tracker.cpp:104
constexpr int kMaxWin = 31;        // LK window sizes supported by the warp-per-feature kernel

struct LevelDesc {
    int w, h, pitch;       // pitch in bytes, multiple of 16
    int64_t offset;        // byte offset of the level inside one stream's pyramid block
};

struct PyrDesc {
    int n_levels;          // levels actually built (<= maxLevel+1)
    LevelDesc lv[kMaxLevels];
    int64_t stream_bytes;  // stride between two streams' pyramid blocks
};

// Camera::liftProjective parameters (f64), see camera.cuh
struct CameraDev {
    int model;
    double inv_K11, inv_K13, inv_K22, inv_K23;
    double d[4];
    double xi;
    double poly[5];
    double cx, cy, inv_scale, aff_c, aff_d, aff_e;
    int no_distortion;
    int kb_npow;
};

__host__ __device__ __forceinline__ int reflect101(int i, int n) {
    // BORDER_REFLECT_101 for -n < i < 2n-1, then clamped so any address stays legal
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
    if (i < 0) i = 0;
    if (i >= n) i = n - 1;
    return i;
}

__device__ __forceinline__ int cv_round(float v) { return __float2int_rn(v); }

}  // namespace vs

#define VS_CUDA_TRY(expr)                                                                     \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) {                                                              \
            vs::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return VS_ERR_CUDA;                                                               \
        }                                                                                     \
    } while (0)
