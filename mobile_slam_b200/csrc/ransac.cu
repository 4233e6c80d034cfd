// Launchers of the device F-matrix RANSAC (algorithm and device code: ransac.cuh).
#include "tracker_state.h"
#include "ransac.cuh"

namespace vs {

__global__ void __launch_bounds__(kRansacThreads) ransac_kernel(GlueParams g, unsigned int seed) {
    pdl_entry();
    extern __shared__ __align__(16) uint8_t dsm[];
    ransac_body(g, seed, blockIdx.x, threadIdx.x, dsm);
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
