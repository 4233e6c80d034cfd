// vs_clahe_no_alloc — CLAHE(3.0, 8x8) on device images for callers that cannot allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
// It is the same two kernels as the frame path / vs_stage_clahe (clahe.cu: tile LUTs, then the bilinear
// blend of the four neighbouring LUTs), so the output is bit-identical to theirs and to cv::CLAHE::apply.
//
// The one piece of scratch CLAHE needs — the 64 tile LUTs of every image, 16 KB per image, which a tracker handle keeps
// in cudaMalloc'ed memory — lives here in a file-scope `static __device__` array sized for kNoAllocMaxImages images.
// Consequences, all documented in include/vs_tracker.h:
//   * every GPU that loads the library holds the array (1 MiB) until the process exits;
//   * ALL calls on a device share it: two calls that overlap on the GPU (different streams, or the same graph launched
//     twice concurrently) race on the LUTs.  Calls issued to one stream are ordered and safe;
//   * it keeps the previous call's contents.  Nothing has to start at zero: the LUT kernel writes all 256 entries of
//     every (image, tile) LUT the blend kernel will read before that kernel starts (stream order), and entries of images
//     beyond n_images are never read — so nothing is cleared.
// In host code `g_luts` is only a host-side stand-in for the symbol, not a device address: its address for the CURRENT
// device comes from cudaGetSymbolAddress, asked once per device and cached.
// The entry point calls no cudaMalloc / cudaFree / allocating library routine, creates no stream or event, and does not
// synchronise: it enqueues two kernels on the caller's stream and returns.
#include <mutex>

#include "../../include/vs_tracker.h"
#include "kernels.h"

namespace {

constexpr int kNoAllocMaxImages = VS_CLAHE_NO_ALLOC_MAX_IMAGES;
constexpr int kLutBytesPerImage = vs::kClaheTiles * vs::kClaheTiles * 256;      // 64 tile LUTs of 256 bytes
constexpr int kMaxDevices = 64;

static __device__ __align__(256) uint8_t g_luts[(size_t)kNoAllocMaxImages * kLutBytesPerImage];

std::mutex g_mu;
uint8_t* g_luts_addr[kMaxDevices] = {};      // device address of g_luts on device d (0 = not asked yet)

// device address of the scratch array on the current device (set-up: once per device, with that device current)
int scratch_address(uint8_t** out) {
    int dev = -1;
    VS_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) { vs::set_error("vs_clahe_no_alloc: device ordinal %d not supported", dev); return VS_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_luts_addr[dev]) {
        void* p = nullptr;
        VS_CUDA_TRY(cudaGetSymbolAddress(&p, g_luts));
        if (!p) { vs::set_error("vs_clahe_no_alloc: cudaGetSymbolAddress returned a null address on device %d", dev); return VS_ERR_CUDA; }
        g_luts_addr[dev] = static_cast<uint8_t*>(p);
    }
    *out = g_luts_addr[dev];
    return VS_OK;
}

}  // namespace

extern "C" int vs_clahe_no_alloc(const uint8_t* d_src, int32_t src_pitch, int64_t src_image_stride, uint8_t* d_dst,
                                 int32_t dst_pitch, int64_t dst_image_stride, int32_t width, int32_t height,
                                 int32_t n_images, void* cuda_stream) {
    using namespace vs;
    if (!d_src || !d_dst) { set_error("vs_clahe_no_alloc: null image pointer"); return VS_ERR_INVALID; }
    if (n_images < 1 || n_images > kNoAllocMaxImages) {
        set_error("vs_clahe_no_alloc: n_images %d outside 1 .. %d (VS_CLAHE_NO_ALLOC_MAX_IMAGES: the static scratch holds the tile LUTs of that many "
                  "images; split the batch into calls on one stream)", n_images, kNoAllocMaxImages);
        return VS_ERR_CAPACITY;
    }
    if (width < 32 || height < 32 || width > 16384 || height > 16384) {
        set_error("vs_clahe_no_alloc: image size %d x %d outside 32 .. 16384 (the limits of vs_tracker_create)", width, height);
        return VS_ERR_INVALID;
    }
    if (src_pitch < width || dst_pitch < width || (src_pitch & 15) || (dst_pitch & 15) || (src_image_stride & 15) || (dst_image_stride & 15) ||
        (reinterpret_cast<uintptr_t>(d_src) & 15) || (reinterpret_cast<uintptr_t>(d_dst) & 15)) {
        set_error("vs_clahe_no_alloc: pointers, pitches and image strides must be multiples of 16 bytes and pitches >= width "
                  "(the kernels use 64-bit row accesses)");
        return VS_ERR_INVALID;
    }
    if (n_images > 1 && (src_image_stride < (int64_t)src_pitch * height || dst_image_stride < (int64_t)dst_pitch * height)) {
        set_error("vs_clahe_no_alloc: image stride below pitch * height");
        return VS_ERR_INVALID;
    }
    uint8_t* luts = nullptr;
    const int rc = scratch_address(&luts);
    if (rc != VS_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    // plain launches: the first kernel must not start before the caller's preceding work on the stream has finished
    const bool pdl_saved = tl_pdl_launch;
    tl_pdl_launch = false;
    launch_clahe(d_src, src_pitch, src_image_stride, d_dst, dst_pitch, dst_image_stride, width, height, n_images, luts, st, nullptr);
    tl_pdl_launch = pdl_saved;
    VS_CUDA_TRY(cudaPeekAtLastError());
    return VS_OK;
}
