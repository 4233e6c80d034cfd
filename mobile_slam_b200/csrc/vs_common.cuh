// Shared device/host definitions for the vs_tracker kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <map>
#include <mutex>
#include <utility>

namespace vs {

constexpr int kMaxLevels = 6;      // maxLevel <= 5  (VIOEngine::setTrackingParams, vio_engine.cpp:369)
constexpr int kClaheTiles = 8;     // createCLAHE(3.0, Size(8,8)), feature_tracker.cpp:104
constexpr int kMaxWin = 31;        // LK window sizes supported by the warp-per-feature kernel

struct LevelDesc {
    int w, h, pitch;       // pitch in bytes, multiple of 16
    int64_t offset;        // byte offset of the level inside one stream's pyramid block
};

// Every level from kPyrFirstPaddedLevel up is stored with a REFLECT_101 border of kPyrPadX columns / kPyrPadY rows
// (cv::buildOpticalFlowPyramid stores its levels padded too): LevelDesc::offset addresses pixel (0, 0), rows -kPyrPadY ..
// h + kPyrPadY - 1 and columns -kPyrPadX .. pitch - kPyrPadX - 1 exist.  The LK kernel's 48 x 32-byte tiles (16-byte chunks,
// origin up to 48 columns / 26 rows outside the level) then need no border handling, where on the coarse levels most tiles
// cross the border.  kPyrFirstPaddedLevel = 1 (level 0 unpadded: it holds half of the border bytes while only ~12 % of its
// tiles touch the border, LK assembles those chunks from reflected bytes) was measured too: border pass 0.102 -> 0.050 ms
// but LK 0.591 -> 0.624 ms per C2 step and +7 us single-stream latency (profiles/r2_experiments.json) — a wash, so every
// level is padded and the fast LK kernel carries no border path.
constexpr int kPyrPadX = 48;       // multiple of 16
constexpr int kPyrPadY = 32;
constexpr int kPyrFirstPaddedLevel = 0;

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

// Programmatic dependent launch.  A frame is a chain of ~10 short kernels on one stream; launched with launch_pdl()
// a kernel may be scheduled onto the SMs while its predecessor drains, which removes the launch gap between them.
// Every kernel starts with pdl_entry(): it blocks until the predecessor grid has completed and its writes are
// visible (so nothing is read or written early), then lets its own successor be scheduled.  Without the launch
// attribute both instructions are no-ops.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_entry() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// Opt a kernel in to more than 48 KB of dynamic shared memory.  The attribute is per device, and one process may
// drive several GPUs (vs::ShardedTracker: one thread + handle per GPU), so what was set is remembered per
// (kernel, device).
inline cudaError_t ensure_dyn_smem(const void* kernel, int bytes) {
    if (bytes <= 48 * 1024) return cudaSuccess;
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, int> done;
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lock(mu);
    int& have = done[std::make_pair(kernel, dev)];
    if (have < bytes) {
        const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        if (e != cudaSuccess) return e;          // the following launch fails too and cudaGetLastError reports it
        have = bytes;
    }
    return cudaSuccess;
}

// Set by the frame enqueue per handle: on for few-stream (latency) handles.  For big stream batches the kernels are
// many waves long, the gaps are negligible, and early-scheduled successors measurably slow the step down (B200:
// 2.43 -> 2.71 ms per 512-stream C2 step), so there the launches stay plain.
inline thread_local bool tl_pdl_launch = false;
constexpr int kPdlMaxStreams = 16;    // measured: +3.5 % at 16 streams, -1.5 % at 64, -10 % at 512 (C1/C2); VS_PDL_MAX_STREAMS overrides

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = tl_pdl_launch ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

}  // namespace vs

#define VS_CUDA_TRY(expr)                                                                     \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) {                                                              \
            vs::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return VS_ERR_CUDA;                                                               \
        }                                                                                     \
    } while (0)
