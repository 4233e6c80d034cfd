// Handle state of the C ABI (include/vs_tracker.h) shared by tracker.cu and ransac.cu.
#pragma once
#include <cstdint>
#include <vector>

#include "../../include/vs_tracker.h"
#include "glue.cuh"

namespace vs {

// kernel slots of the per-kernel CUDA-event profile (vs_tracker_get_profile)
enum ProfSlot { PK_INPUT = 0, PK_CLAHE_LUT, PK_CLAHE_APPLY, PK_PYR_DOWN, PK_PYR_PAD, PK_LK, PK_LK_BACK, PK_POST_LK, PK_RANSAC,
                PK_SET_MASK, PK_SHI_TOMASI, PK_SELECT_FINISH, PK_COUNT };

struct Profiler {
    bool on = false;
    std::vector<cudaEvent_t> ev;      // event pool, grown on demand
    std::vector<int> slot;            // slot[i] = kernel that ran between ev[i] and ev[i+1]
    int used = 0;
    double total_ms[PK_COUNT] = {};
    int launches[PK_COUNT] = {};
};

struct DeviceArena {
    // Every block carries kSlack bytes after its payload: vector loads of the kernels may read (never write) a few bytes
    // past the last element.  With VS_DEBUG_GUARD=1 in the environment at handle creation the block is additionally
    // bracketed by two kGuard-byte bands of 0xA5 that nothing may write; check() counts the bytes that changed
    // (compute-sanitizer is not available on every box — this is the repo's own out-of-bounds-write detector).
    static constexpr size_t kSlack = 256, kGuard = 4096;
    struct Block { void* base; size_t payload; };
    std::vector<Block> blocks;
    bool guarded = false;
    template <class T>
    cudaError_t alloc(T** p, size_t count, bool zero = true) {
        const size_t payload = (count * sizeof(T) + 15) & ~(size_t)15;
        const size_t lead = guarded ? kGuard : 0, tail = guarded ? kGuard : kSlack;
        void* q = nullptr;
        cudaError_t e = cudaMalloc(&q, lead + payload + tail);
        if (e != cudaSuccess) return e;
        blocks.push_back({q, payload});
        uint8_t* b = static_cast<uint8_t*>(q);
        if (guarded) {
            e = cudaMemset(b, 0xA5, lead);
            if (e == cudaSuccess) e = cudaMemset(b + lead + payload, 0xA5, tail);
            if (e != cudaSuccess) return e;
        }
        if (zero) {
            e = cudaMemset(b + lead, 0, payload + (guarded ? 0 : tail));
            if (e != cudaSuccess) return e;
        }
        *p = reinterpret_cast<T*>(b + lead);
        return cudaSuccess;
    }
    // number of guard bytes that no longer hold 0xA5 (-1: guards were not enabled, -2: CUDA error)
    int64_t check() const {
        if (!guarded) return -1;
        std::vector<uint8_t> h(kGuard);
        int64_t bad = 0;
        for (const Block& bk : blocks) {
            const uint8_t* b = static_cast<const uint8_t*>(bk.base);
            for (int side = 0; side < 2; ++side) {
                if (cudaMemcpy(h.data(), b + (side ? kGuard + bk.payload : 0), kGuard, cudaMemcpyDeviceToHost) != cudaSuccess) return -2;
                for (uint8_t v : h) bad += v != 0xA5;
            }
        }
        return bad;
    }
    void release() {
        for (const Block& b : blocks) cudaFree(b.base);
        blocks.clear();
    }
};

}  // namespace vs

struct vs_tracker {
    vs_tracker_config cfg{};
    int B = 0, W = 0, H = 0, cap = 0;
    int raw_pitch = 0;
    int64_t raw_stride = 0;
    vs::PyrDesc pd{};                 // geometry of every possible level (kMaxLevels)
    int built[2] = {0, 0};        // levels built in each pyramid buffer
    int cur = 0;                  // index of the reference (previous-frame) pyramid
    bool has_prev = false;
    int n_bound = 0;              // upper bound of per-stream feature counts (LK grid size)
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_ms = 0.f;
    int64_t launches = 0;
    vs::DeviceArena arena;
    uint8_t* d_raw = nullptr;
    uint8_t* d_pyr[2] = {nullptr, nullptr};
    uint8_t* d_luts = nullptr;
    uint8_t* d_base_mask = nullptr;
    int* d_disc_hw = nullptr;
    int* d_fish_hw = nullptr;     // row half widths of the auto fisheye circle
    int base_mode = 0, fcx = 0, fcy = 0, fr = 0;   // StParams::base_mode and the circle
    double* d_cur_time = nullptr;
    int* d_lk_counters = nullptr; // persistent-warp LK work queue heads
    vs::LkTmaMaps* lk_tma = nullptr;   // tensor maps of d_pyr[0..1] (VS_LK_TMA=1 only), host memory
    int lk_seq = 0;
    float* d_eig = nullptr;       // lazily allocated dense eig map (stage API only)
    vs::GlueParams g{};
    // pinned host mirrors
    uint8_t* h_block = nullptr;   // results: hdr | ids | cnt | pts | un | vel
    size_t h_block_bytes = 0;
    uint8_t* d_rblock = nullptr;
    const uint8_t* last_results = nullptr;   // block the last frame's kernels wrote (device block, or the pinned slot of a small handle)
    // timestamps are read in place from pinned host memory by select_finish: a ring of slots, each guarded by an event
    // recorded after the frame that reads it, so back-to-back unsynchronised calls never overwrite a slot still in use
    static constexpr int kTimeRing = 8;
    double* h_time = nullptr;
    cudaEvent_t ev_time[kTimeRing] = {};
    int64_t time_seq = 0;
    // RANSAC callback hop
    vs_reject_fn reject_fn = nullptr;
    void* reject_user = nullptr;
    int* h_wn = nullptr;
    int* h_flag = nullptr;
    float2* h_lcur = nullptr;
    float2* h_lnext = nullptr;
    float2* h_wnext = nullptr;
    uint8_t* h_rstatus = nullptr;
    std::vector<int> disc_hw_host;
    int disc_radius = -1;
    vs::Profiler prof;
    // pipeline lanes (cfg.pipeline_lanes > 1): this handle then owns no device buffers of its own, only `stream`
    // (fork/join point) and the lanes; lane l tracks streams [lane_s0[l], lane_s0[l+1])
    std::vector<vs_tracker*> lanes;
    std::vector<int> lane_s0;
    cudaEvent_t ev_done = nullptr;    // recorded on `stream` after the lane's last kernel (join)
    cudaEvent_t ev_fetch[2] = {nullptr, nullptr};   // result slot k downloaded
    int64_t submit_seq = 0, wait_seq = 0;            // vs_tracker_submit_batch / vs_tracker_wait
};

