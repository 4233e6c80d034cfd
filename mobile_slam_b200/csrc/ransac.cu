// Launchers of the device F-matrix RANSAC (algorithm and device code: ransac.cuh).
#include "tracker_state.h"
#include "ransac.cuh"

namespace vs {

__global__ void __launch_bounds__(kRansacThreads) ransac_kernel(GlueParams g, unsigned int seed) {
    pdl_entry();
    extern __shared__ __align__(16) uint8_t dsm[];
    ransac_body(g, seed, blockIdx.x, threadIdx.x, dsm);
}

struct FMat { double f[9]; };
// teacher-forced final marking (vs_stage_ransac_mark): flags for a GIVEN model, points as uploaded (no centring)
__global__ void __launch_bounds__(kRansacThreads) ransac_mark_kernel(GlueParams g, FMat F, int keep_base) {
    ransac_mark(g, 0, g.wn[0], F.f, g.w_lcur, g.w_lnext, keep_base != 0, threadIdx.x, kRansacThreads);
}

static int ransac_smem(int cap) { return cap * 16; }

int launch_ransac(vs_tracker* t) {
    const int smem = ransac_smem(t->cap);
    ensure_dyn_smem(reinterpret_cast<const void*>(ransac_kernel), smem);
    launch_pdl(ransac_kernel, dim3(t->B), dim3(kRansacThreads), smem, t->stream, t->g, 12345u /* = ransac_seed() in tracker.cu */);
    t->launches += 1;
    return VS_OK;
}

}  // namespace vs

extern "C" int vs_stage_ransac(vs_tracker* t, const float* a_xy, const float* b_xy, int32_t n, uint8_t* status) {
    using namespace vs;
    if (!t || !a_xy || !b_xy || !status || n < 8) { set_error("vs_stage_ransac: bad argument (need n >= 8)"); return VS_ERR_INVALID; }
    if (!t->lanes.empty()) t = t->lanes[0];
    if (n > t->cap) { set_error("n %d above capacity %d", n, t->cap); return VS_ERR_CAPACITY; }
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    GlueParams& g = t->g;
    g.f_threshold = t->cfg.f_threshold;
    g.f_edge_factor = 0.0;                       // the stage takes lifted points only (no pixel positions)
    const int one = 1;
    VS_CUDA_TRY(cudaMemcpyAsync(g.w_lcur, a_xy, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(g.w_lnext, b_xy, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(g.wn, &n, sizeof(int), cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(g.ransac_flag, &one, sizeof(int), cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(g.w_rstatus, 1, (size_t)n, t->stream));
    GlueParams g1 = g;
    g1.n_streams = 1;
    const int smem = ransac_smem(t->cap);
    ensure_dyn_smem(reinterpret_cast<const void*>(ransac_kernel), smem);
    ransac_kernel<<<1, kRansacThreads, smem, t->stream>>>(g1, 12345u);
    t->launches += 1;
    VS_CUDA_TRY(cudaMemcpyAsync(status, g.w_rstatus, (size_t)n, cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(g.n, 0, sizeof(int) * t->B, t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(g.ransac_flag, 0, sizeof(int), t->stream));
    VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
    VS_CUDA_TRY(cudaGetLastError());
    t->has_prev = false;
    return VS_OK;
}

// Given a fundamental matrix (row-major, x2^T F x1 = 0; e.g. cv::findFundamentalMat's) the flags the device assigns:
// base_status == NULL: inlier test (both epipolar distances <= f_threshold) + edge recovery; base_status != NULL: the
// teacher's flags + only the edge-recovery rule of feature_tracker.cpp:236-285 (edge_factor <= 0 switches it off).
extern "C" int vs_stage_ransac_mark(vs_tracker* t, const double* f9, const float* a_xy, const float* b_xy, const float* next_px_xy,
                                    int32_t n, double edge_factor, const uint8_t* base_status, uint8_t* status) {
    using namespace vs;
    if (!t || !f9 || !a_xy || !b_xy || !next_px_xy || !status || n < 1) { set_error("vs_stage_ransac_mark: bad argument"); return VS_ERR_INVALID; }
    if (!t->lanes.empty()) t = t->lanes[0];
    if (n > t->cap) { set_error("n %d above capacity %d", n, t->cap); return VS_ERR_CAPACITY; }
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    GlueParams g1 = t->g;
    g1.n_streams = 1;
    g1.f_threshold = t->cfg.f_threshold;
    g1.f_edge_factor = edge_factor > 0.0 ? edge_factor : 0.0;
    VS_CUDA_TRY(cudaMemcpyAsync(g1.w_lcur, a_xy, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(g1.w_lnext, b_xy, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(g1.w_next, next_px_xy, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(g1.wn, &n, sizeof(int), cudaMemcpyHostToDevice, t->stream));
    if (base_status) VS_CUDA_TRY(cudaMemcpyAsync(g1.w_rstatus, base_status, (size_t)n, cudaMemcpyHostToDevice, t->stream));
    FMat F;
    for (int i = 0; i < 9; ++i) F.f[i] = f9[i];
    ransac_mark_kernel<<<1, kRansacThreads, 0, t->stream>>>(g1, F, base_status ? 1 : 0);
    t->launches += 1;
    VS_CUDA_TRY(cudaMemcpyAsync(status, g1.w_rstatus, (size_t)n, cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(t->g.n, 0, sizeof(int) * t->B, t->stream));
    VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
    VS_CUDA_TRY(cudaGetLastError());
    t->has_prev = false;
    return VS_OK;
}
