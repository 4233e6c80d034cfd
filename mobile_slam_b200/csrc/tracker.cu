// Host layer of the B200 visual front-end: the C ABI of include/vs_tracker.h over CUDA streams.
// Mirrors frontend::FeatureTracker (include/frontend/feature_tracker.h:25-76 of the reference):
// one handle = n_streams independent trackers whose per-frame work is batched into the same launches.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "tracker_state.h"

namespace vs {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// OpenCV's filled-circle row half-widths (cv::circle, thickness -1, LINE_8): integer midpoint
// rasteriser, SURVEY.md S5.  hw[|dy|] = -1 when the row is not painted.
static std::vector<int> circle_half_widths(int radius) {
    std::vector<int> hw(radius + 1, -1);
    int dx = radius, dy = 0, err = 0, plus = 1, minus = (radius << 1) - 1;
    while (dx >= dy) {
        if (dy <= radius && hw[dy] < dx) hw[dy] = dx;
        if (dx <= radius && hw[dx] < dy) hw[dx] = dy;
        dy++;
        err += plus;
        plus += 2;
        const int mask = (err <= 0) - 1;
        err -= minus & mask;
        dx += mask;
        minus -= mask & 2;
    }
    return hw;
}

static inline int round_up(int v, int a) { return (v + a - 1) / a * a; }

}  // namespace vs

using namespace vs;

namespace vs {

static void plan_pyramid(vs_tracker* t) {
    PyrDesc& pd = t->pd;
    int w = t->W, h = t->H;
    int64_t off = 0;
    pd.n_levels = kMaxLevels;
    for (int l = 0; l < kMaxLevels; ++l) {
        pd.lv[l].w = w;
        pd.lv[l].h = h;
        const int px = l >= kPyrFirstPaddedLevel ? kPyrPadX : 0, py = l >= kPyrFirstPaddedLevel ? kPyrPadY : 0;
        pd.lv[l].pitch = round_up(w + 2 * px, 16);
        pd.lv[l].offset = off + (int64_t)py * pd.lv[l].pitch + px;                  // pixel (0, 0) inside the padded block
        off += (int64_t)round_up(pd.lv[l].pitch * (h + 2 * py), 256);
        w = (w + 1) / 2;
        h = (h + 1) / 2;
        if (w < 1) w = 1;
        if (h < 1) h = 1;
    }
    pd.stream_bytes = off;
}

// levels cv::buildOpticalFlowPyramid would produce for (winSize, maxLevel)
static int levels_to_build(const vs_tracker* t) {
    int n = 1;
    int w = t->W, h = t->H;
    for (int l = 0; l < t->cfg.lk_pyramid_levels && l + 1 < kMaxLevels; ++l) {
        const int nw = (w + 1) / 2, nh = (h + 1) / 2;
        if (nw <= t->cfg.lk_window_size || nh <= t->cfg.lk_window_size) break;
        w = nw;
        h = nh;
        ++n;
    }
    return n;
}

static void fill_camera(const vs_camera& c, CameraDev& d) {
    memset(&d, 0, sizeof(d));
    d.model = c.model_type;
    if (c.model_type != VS_CAM_SCARAMUZZA) {
        d.inv_K11 = 1.0 / c.fx;
        d.inv_K13 = -c.cx / c.fx;
        d.inv_K22 = 1.0 / c.fy;
        d.inv_K23 = -c.cy / c.fy;
    }
    for (int i = 0; i < 4; ++i) d.d[i] = c.d[i];
    d.xi = c.xi;
    for (int i = 0; i < 5; ++i) d.poly[i] = c.poly[i];
    d.cx = c.cx;
    d.cy = c.cy;
    d.aff_c = c.aff_c;
    d.aff_d = c.aff_d;
    d.aff_e = c.aff_e;
    const double det = c.aff_c - c.aff_d * c.aff_e;
    d.inv_scale = det != 0.0 ? 1.0 / det : 0.0;
    d.no_distortion = (c.d[0] == 0.0 && c.d[1] == 0.0 && c.d[2] == 0.0 && c.d[3] == 0.0) ? 1 : 0;
    int npow = 9;   // EquidistantCamera::backprojectSymmetric, EquidistantCamera.cc:585-597
    if (c.d[3] == 0.0) npow -= 2;
    if (c.d[2] == 0.0) npow -= 2;
    if (c.d[1] == 0.0) npow -= 2;
    if (c.d[0] == 0.0) npow -= 2;
    d.kb_npow = npow;
}

static int upload_disc(vs_tracker* t) {
    const int r = t->cfg.min_dist;
    if (r == t->disc_radius) return VS_OK;
    t->disc_hw_host = circle_half_widths(r);
    VS_CUDA_TRY(cudaMemcpyAsync(t->d_disc_hw, t->disc_hw_host.data(), sizeof(int) * (r + 1), cudaMemcpyHostToDevice,
                                t->stream));
    VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
    t->disc_radius = r;
    return VS_OK;
}

// The mask setMask starts from (feature_tracker.cpp:40-60): all 255, or with `fisheye` the mask image if one is given,
// else the auto-generated circle.  The reference builds its fisheye mask lazily on the first frame that has the flag
// set and keeps it; here it is (re)built whenever the flag or the image changes (vs_tracker_set_frontend_flags).
static int build_base_mask(vs_tracker* t, int fisheye, const uint8_t* mask_image, int mask_stride) {
    GlueParams& g = t->g;
    std::vector<uint8_t> m((size_t)g.mask_stride, 0);
    const int W = t->W, H = t->H, mp = g.mask_pitch;
    if (fisheye && mask_image) {
        const int st = mask_stride > 0 ? mask_stride : W;
        for (int y = 0; y < H; ++y) memcpy(&m[(size_t)y * mp], mask_image + (size_t)y * st, W);
        t->base_mode = 2;
    } else if (fisheye) {
        const double cx = W / 2.0, cy = H / 2.0;
        double rad = cx;
        if (cy < rad) rad = cy;
        if (W - cx < rad) rad = W - cx;
        if (H - cy < rad) rad = H - cy;
        rad *= 0.95;
        const int icx = (int)cx, icy = (int)cy, r = (int)rad;
        const std::vector<int> hw = circle_half_widths(r);
        t->base_mode = 1;
        t->fcx = icx;
        t->fcy = icy;
        t->fr = r;
        VS_CUDA_TRY(cudaMemcpy(t->d_fish_hw, hw.data(), sizeof(int) * (r + 1), cudaMemcpyHostToDevice));
        for (int dy = -r; dy <= r; ++dy) {
            const int y = icy + dy, w = hw[dy < 0 ? -dy : dy];
            if (y < 0 || y >= H || w < 0) continue;
            int xa = icx - w, xb = icx + w;
            if (xa < 0) xa = 0;
            if (xb > W - 1) xb = W - 1;
            for (int x = xa; x <= xb; ++x) m[(size_t)y * mp + x] = 255;
        }
    } else {
        t->base_mode = 0;
        for (int y = 0; y < H; ++y) memset(&m[(size_t)y * mp], 255, W);
    }
    VS_CUDA_TRY(cudaMemcpy(t->d_base_mask, m.data(), m.size(), cudaMemcpyHostToDevice));
    t->cfg.fisheye = fisheye ? 1 : 0;
    return VS_OK;
}

static void sync_glue_params(vs_tracker* t) {
    GlueParams& g = t->g;
    g.max_cnt = t->cfg.max_cnt;
    g.min_dist = t->cfg.min_dist;
    g.ransac_enabled = t->cfg.ransac_mode != VS_RANSAC_OFF;
    g.focal_length = t->cfg.focal_length;
    g.f_threshold = t->cfg.f_threshold;
    g.f_edge_factor = t->cfg.f_threshold_edge_factor;
    g.fb_check = t->cfg.fb_check;
    g.fb_threshold = (float)t->cfg.fb_threshold;
}


// results block: hdr | ids | cnt | pts | un | vel, in device memory (d_rblock, copied out after the frame) or — for
// latency-sized handles — directly in a pinned host slot (the kernels' few KB of result stores go over PCIe as posted
// writes, which saves the separate D2H copy and its DMA latency at the end of every frame)
constexpr int kDirectResultsMaxStreams = 16;
static void point_results(vs_tracker* t, uint8_t* base) {
    GlueParams& g = t->g;
    const int64_t B = t->B, C = t->cap;
    uint8_t* p = base;
    g.r_hdr = reinterpret_cast<int*>(p); p += B * kHdrInts * 4;
    g.r_ids = reinterpret_cast<int*>(p); p += B * C * 4;
    g.r_cnt = reinterpret_cast<int*>(p); p += B * C * 4;
    g.r_pts = reinterpret_cast<float2*>(p); p += B * C * 8;
    g.r_un = reinterpret_cast<float2*>(p); p += B * C * 8;
    g.r_vel = reinterpret_cast<float2*>(p);
}
static inline bool direct_results(const vs_tracker* t) { return t->B <= kDirectResultsMaxStreams; }

// ---- per-kernel CUDA-event profile (vs_tracker_set_profiling) ----
static const char* const kProfNames[PK_COUNT] = {
    "input_copy", "clahe_lut", "clahe_apply", "pyr_down", "pyr_pad", "lk", "lk_backward", "post_lk", "ransac",
    "set_mask", "shi_tomasi", "select_finish"};

static void prof_begin(vs_tracker* t) {
    Profiler& p = t->prof;
    if (!p.on) return;
    p.used = 0;
    p.slot.clear();
    if (p.ev.empty()) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        p.ev.push_back(e);
    }
    cudaEventRecord(p.ev[0], t->stream);
    p.used = 1;
}

// the kernels launched since the previous tick belong to `slot`
static void prof_tick(vs_tracker* t, int slot) {
    Profiler& p = t->prof;
    if (!p.on) return;
    if ((int)p.ev.size() <= p.used) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        p.ev.push_back(e);
    }
    cudaEventRecord(p.ev[p.used], t->stream);
    p.slot.push_back(slot);
    ++p.used;
}

static void prof_end(vs_tracker* t) {
    Profiler& p = t->prof;
    if (!p.on || p.used < 2) return;
    cudaEventSynchronize(p.ev[p.used - 1]);
    for (int i = 0; i + 1 < p.used; ++i) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, p.ev[i], p.ev[i + 1]) == cudaSuccess) {
            p.total_ms[p.slot[i]] += ms;
            p.launches[p.slot[i]] += 1;
        }
    }
}

}  // namespace vs

// =================================================================================================
// C ABI
// =================================================================================================
extern "C" {

const char* vs_last_error(void) { return vs::g_err; }
int vs_abi_version(void) { return VS_ABI_VERSION; }

void vs_tracker_default_config(vs_tracker_config* c) {
    if (!c) return;
    memset(c, 0, sizeof(*c));
    c->struct_size = (int32_t)sizeof(*c);
    c->n_streams = 1;
    c->width = 752;                 // include/utility/config.h:18-30 defaults
    c->height = 480;
    c->focal_length = 460.0;
    c->max_cnt = 150;
    c->min_dist = 30;
    c->f_threshold = 1.0;
    c->f_threshold_edge_factor = 0.0;
    c->equalize = 1;
    c->fisheye = 0;
    c->lk_window_size = 21;
    c->lk_pyramid_levels = 3;
    c->lk_iterations = 30;
    c->lk_eps = 0.01;
    c->fb_check = 0;
    c->fb_threshold = 0.5;
    c->ransac_mode = VS_RANSAC_DEVICE;
    c->pipeline_lanes = 0;
    c->camera.model_type = VS_CAM_PINHOLE;
    c->camera.width = 752;
    c->camera.height = 480;
    c->camera.fx = c->camera.fy = 460.0;
    c->camera.cx = 376.0;
    c->camera.cy = 240.0;
    c->camera.aff_c = 1.0;
}

static const int kMaxDiscRadius = 1024;

int vs_tracker_create(const vs_tracker_config* cfg, vs_tracker_config* effective, vs_tracker** out) {
    if (!cfg || !out) { set_error("vs_tracker_create: null argument"); return VS_ERR_INVALID; }
    *out = nullptr;
    if (cfg->struct_size != (int32_t)sizeof(vs_tracker_config)) {
        set_error("vs_tracker_create: struct_size %d != %d (ABI mismatch)", cfg->struct_size, (int)sizeof(vs_tracker_config));
        return VS_ERR_INVALID;
    }
    if (cfg->n_streams < 1 || cfg->width < 32 || cfg->height < 32 || cfg->width > 16384 || cfg->height > 16384) {
        set_error("vs_tracker_create: bad n_streams/width/height");
        return VS_ERR_INVALID;
    }
    if (cfg->max_cnt < 1 || cfg->min_dist < 1 || cfg->min_dist > kMaxDiscRadius) {
        set_error("vs_tracker_create: bad max_cnt/min_dist");
        return VS_ERR_INVALID;
    }
    if (cfg->lk_window_size < 3 || cfg->lk_window_size > kMaxWin || (cfg->lk_window_size & 1) == 0) {
        set_error("vs_tracker_create: lk_window_size must be odd and in [3,%d]", kMaxWin);
        return VS_ERR_INVALID;
    }
    if (cfg->lk_pyramid_levels < 0 || cfg->lk_pyramid_levels > kMaxLevels - 1) {
        set_error("vs_tracker_create: lk_pyramid_levels must be in [0,%d]", kMaxLevels - 1);
        return VS_ERR_INVALID;
    }
    if (cfg->camera.model_type < 0 || cfg->camera.model_type > 3) {
        set_error("vs_tracker_create: unknown camera model %d", cfg->camera.model_type);
        return VS_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        set_error("vs_tracker_create: no CUDA device (this library has no CPU fallback)");
        return VS_ERR_CUDA;
    }
    if (cfg->device < 0 || cfg->device >= ndev) { set_error("vs_tracker_create: bad device %d", cfg->device); return VS_ERR_INVALID; }
    VS_CUDA_TRY(cudaSetDevice(cfg->device));

    vs_tracker* t = new (std::nothrow) vs_tracker();
    if (!t) { set_error("out of host memory"); return VS_ERR_NOMEM; }
    t->cfg = *cfg;
    t->cfg.fisheye_mask_image = nullptr;
    t->B = cfg->n_streams;
    t->W = cfg->width;
    t->H = cfg->height;
    int cap = cfg->capacity > 0 ? cfg->capacity : (cfg->max_cnt > 256 ? cfg->max_cnt : 256);
    if (cap < cfg->max_cnt) cap = cfg->max_cnt;
    cap = round_up(cap, 32);
    if (cap > 4096) { delete t; set_error("capacity above 4096 features per stream is not supported"); return VS_ERR_CAPACITY; }
    {
        // the per-stream glue kernels keep a stream's whole feature set in shared memory: refuse a capacity they cannot
        // hold here, instead of failing at the first track call with a launch error
        int smem_optin = 0;
        cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, cfg->device);
        const int need = glue_max_smem_bytes(cap) + 2048;      // + the kernels' static shared memory
        if (need > smem_optin) {
            delete t;
            set_error("capacity %d needs %d B of shared memory per CTA, the device allows %d", cap, need, smem_optin);
            return VS_ERR_CAPACITY;
        }
    }
    t->cap = cap;
    t->cfg.capacity = cap;

    // ---- pipeline lanes: the handle becomes a fork/join point over sub-handles, one CUDA stream each ----
    int n_lanes = cfg->pipeline_lanes;
    if (n_lanes > cfg->n_streams) n_lanes = cfg->n_streams;
    if (cfg->ransac_mode == VS_RANSAC_CALLBACK) n_lanes = 1;     // the callback hop is a per-handle host sync
    if (n_lanes > 1) {
        t->cfg.pipeline_lanes = n_lanes;
        cudaError_t e = cudaStreamCreateWithFlags(&t->stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreate(&t->ev0);
        if (e == cudaSuccess) e = cudaEventCreate(&t->ev1);
        if (e != cudaSuccess) { set_error("lane fork/join setup: %s", cudaGetErrorString(e)); vs_tracker_destroy(t); return VS_ERR_CUDA; }
        int s0 = 0;
        for (int l = 0; l < n_lanes; ++l) {
            const int nl = (cfg->n_streams - s0) / (n_lanes - l);
            vs_tracker_config lc = *cfg;
            lc.n_streams = nl;
            lc.pipeline_lanes = 1;
            lc.capacity = cap;
            vs_tracker* lane = nullptr;
            const int rc = vs_tracker_create(&lc, nullptr, &lane);
            if (rc != VS_OK) { vs_tracker_destroy(t); return rc; }
            if (cudaEventCreateWithFlags(&lane->ev_done, cudaEventDisableTiming) != cudaSuccess) {
                vs_tracker_destroy(lane); vs_tracker_destroy(t); set_error("lane event"); return VS_ERR_CUDA;
            }
            t->lanes.push_back(lane);
            t->lane_s0.push_back(s0);
            s0 += nl;
        }
        t->lane_s0.push_back(s0);
        if (effective) *effective = t->cfg;
        *out = t;
        return VS_OK;
    }
    t->cfg.pipeline_lanes = 1;
    t->n_bound = cfg->max_cnt;
    plan_pyramid(t);
    t->raw_pitch = round_up(t->W, 16);
    t->raw_stride = (int64_t)round_up(t->raw_pitch * t->H, 256);

#define VS_TRY_CREATE(expr)                                                                      \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess) {                                                                 \
            set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));     \
            vs_tracker_destroy(t);                                                               \
            return _e == cudaErrorMemoryAllocation ? VS_ERR_NOMEM : VS_ERR_CUDA;                 \
        }                                                                                        \
    } while (0)

    VS_TRY_CREATE(cudaStreamCreateWithFlags(&t->stream, cudaStreamNonBlocking));
    VS_TRY_CREATE(cudaEventCreate(&t->ev0));
    VS_TRY_CREATE(cudaEventCreate(&t->ev1));
    const int64_t B = t->B, C = t->cap;
    GlueParams& g = t->g;
    g.n_streams = t->B;
    g.cap = t->cap;
    g.W = t->W;
    g.H = t->H;
    fill_camera(cfg->camera, g.cam);
    sync_glue_params(t);
    DeviceArena& A = t->arena;
    {
        const char* e = getenv("VS_DEBUG_GUARD");
        A.guarded = e && atoi(e) != 0;
    }
    VS_TRY_CREATE(A.alloc(&t->d_raw, (size_t)(B * t->raw_stride)));
    VS_TRY_CREATE(A.alloc(&t->d_pyr[0], (size_t)(B * t->pd.stream_bytes)));
    VS_TRY_CREATE(A.alloc(&t->d_pyr[1], (size_t)(B * t->pd.stream_bytes)));
    VS_TRY_CREATE(A.alloc(&t->d_luts, (size_t)(B * 64 * 256)));
    VS_TRY_CREATE(A.alloc(&t->d_disc_hw, (size_t)kMaxDiscRadius + 1));
    VS_TRY_CREATE(A.alloc(&t->d_fish_hw, (size_t)(t->W > t->H ? t->W : t->H) + 2));
    VS_TRY_CREATE(A.alloc(&t->d_cur_time, (size_t)B));
    VS_TRY_CREATE(A.alloc(&t->d_lk_counters, (size_t)2));
    {   // A/B variant of the LK tile staging (VS_LK_TMA=1): tensor maps over the two pyramid buffers
        const char* e = getenv("VS_LK_TMA");
        if (e && atoi(e) != 0) {
            t->lk_tma = new LkTmaMaps();
            if (lk_build_tma_maps(t->lk_tma, t->d_pyr, t->pd, t->B) != 0) { delete t->lk_tma; t->lk_tma = nullptr; }
        }
    }
    g.cur_time = t->d_cur_time;
    VS_TRY_CREATE(A.alloc(&g.n, B));
    VS_TRY_CREATE(A.alloc(&g.pts, B * C));
    VS_TRY_CREATE(A.alloc(&g.ids, B * C));
    VS_TRY_CREATE(A.alloc(&g.cnt, B * C));
    VS_TRY_CREATE(A.alloc(&g.un, B * C));
    VS_TRY_CREATE(A.alloc(&g.next_id, B));
    VS_TRY_CREATE(A.alloc(&g.prev_time, B));
    VS_TRY_CREATE(A.alloc(&g.lk_next, B * C));
    VS_TRY_CREATE(A.alloc(&g.lk_status, B * C));
    VS_TRY_CREATE(A.alloc(&g.lk_back, B * C));
    VS_TRY_CREATE(A.alloc(&g.lk_back_status, B * C));
    VS_TRY_CREATE(A.alloc(&g.wn, B));
    VS_TRY_CREATE(A.alloc(&g.w_cur, B * C));
    VS_TRY_CREATE(A.alloc(&g.w_next, B * C));
    VS_TRY_CREATE(A.alloc(&g.w_ids, B * C));
    VS_TRY_CREATE(A.alloc(&g.w_cnt, B * C));
    VS_TRY_CREATE(A.alloc(&g.w_un, B * C));
    VS_TRY_CREATE(A.alloc(&g.w_lcur, B * C));
    VS_TRY_CREATE(A.alloc(&g.w_lnext, B * C));
    VS_TRY_CREATE(A.alloc(&g.w_rstatus, B * C));
    VS_TRY_CREATE(A.alloc(&g.ransac_flag, B));
    VS_TRY_CREATE(A.alloc(&g.kn, B));
    VS_TRY_CREATE(A.alloc(&g.k_pts, B * C));
    VS_TRY_CREATE(A.alloc(&g.k_ids, B * C));
    VS_TRY_CREATE(A.alloc(&g.k_cnt, B * C));
    VS_TRY_CREATE(A.alloc(&g.k_un, B * C));
    VS_TRY_CREATE(A.alloc(&g.want, B));
    g.mask_pitch = round_up(t->W, 16);
    g.mask_stride = (int64_t)round_up(g.mask_pitch * t->H, 256);
    VS_TRY_CREATE(A.alloc(&g.mask, (size_t)g.mask_stride));     // stage API only: the frame path never materialises a mask
    VS_TRY_CREATE(A.alloc(&t->d_base_mask, (size_t)g.mask_stride));
    g.base_mask = t->d_base_mask;
    g.disc_hw = t->d_disc_hw;
    // 3x3 NMS is non-strict (v >= neighbours, like OpenCV's eig == dilate(eig)), so plateaus of equal response can make
    // every pixel of a region a candidate: size the list for the whole image, then nothing is ever dropped
    const int cc = round_up(t->W * t->H, 64);
    g.cand_cap = cc;
    VS_TRY_CREATE(A.alloc(&g.cand, (size_t)(B * cc), false));
    VS_TRY_CREATE(A.alloc(&g.cand_count, B));
    VS_TRY_CREATE(A.alloc(&g.max_enc, B));
    {   // the record list of the two-tier Shi-Tomasi variant, only when that variant is selected (A/B runs)
        const char* e = getenv("VS_ST_TWO_TIER");
        if (e && atoi(e) != 0) VS_TRY_CREATE(A.alloc(&g.cand_raw, (size_t)(B * cc), false));
    }
    VS_TRY_CREATE(A.alloc(&g.raw_count, B));
    VS_TRY_CREATE(A.alloc(&g.l_enc, B));
    // results block: hdr | ids | cnt | pts | un | vel
    t->h_block_bytes = (size_t)(B * kHdrInts * 4 + B * C * (4 + 4 + 8 + 8 + 8));
    VS_TRY_CREATE(A.alloc(&t->d_rblock, t->h_block_bytes));
    point_results(t, t->d_rblock);
    VS_TRY_CREATE(cudaMallocHost(&t->h_block, 2 * t->h_block_bytes));      // two result slots (submit / wait pipelining)
    memset(t->h_block, 0, 2 * t->h_block_bytes);
    VS_TRY_CREATE(cudaMallocHost(&t->h_time, vs_tracker::kTimeRing * sizeof(double) * B));
    for (int k = 0; k < vs_tracker::kTimeRing; ++k) VS_TRY_CREATE(cudaEventCreateWithFlags(&t->ev_time[k], cudaEventDisableTiming));
    VS_TRY_CREATE(cudaEventCreateWithFlags(&t->ev_fetch[0], cudaEventDisableTiming));
    VS_TRY_CREATE(cudaEventCreateWithFlags(&t->ev_fetch[1], cudaEventDisableTiming));
    VS_TRY_CREATE(cudaMallocHost(&t->h_wn, sizeof(int) * B));
    VS_TRY_CREATE(cudaMallocHost(&t->h_flag, sizeof(int) * B));
    VS_TRY_CREATE(cudaMallocHost(&t->h_lcur, sizeof(float2) * B * C));
    VS_TRY_CREATE(cudaMallocHost(&t->h_lnext, sizeof(float2) * B * C));
    VS_TRY_CREATE(cudaMallocHost(&t->h_wnext, sizeof(float2) * B * C));
    VS_TRY_CREATE(cudaMallocHost(&t->h_rstatus, (size_t)(B * C)));

    {
        const int rc = build_base_mask(t, cfg->fisheye, cfg->fisheye_mask_image, cfg->fisheye_mask_stride);
        if (rc != VS_OK) { vs_tracker_destroy(t); return rc; }
    }
    {
        std::vector<int> ids((size_t)B, cfg->first_feature_id);
        VS_TRY_CREATE(cudaMemcpy(g.next_id, ids.data(), sizeof(int) * B, cudaMemcpyHostToDevice));
    }
    if (upload_disc(t) != VS_OK) { vs_tracker_destroy(t); return VS_ERR_CUDA; }
    VS_TRY_CREATE(cudaStreamSynchronize(t->stream));
#undef VS_TRY_CREATE
    if (effective) *effective = t->cfg;
    *out = t;
    return VS_OK;
}

void vs_tracker_destroy(vs_tracker* t) {
    if (!t) return;
    cudaSetDevice(t->cfg.device);
    for (vs_tracker* lane : t->lanes) vs_tracker_destroy(lane);
    t->lanes.clear();
    if (t->ev_done) cudaEventDestroy(t->ev_done);
    for (int k = 0; k < 2; ++k) if (t->ev_fetch[k]) cudaEventDestroy(t->ev_fetch[k]);
    for (int k = 0; k < vs_tracker::kTimeRing; ++k) if (t->ev_time[k]) cudaEventDestroy(t->ev_time[k]);
    if (t->stream) cudaStreamSynchronize(t->stream);
    t->arena.release();
    if (t->d_eig) cudaFree(t->d_eig);
    delete t->lk_tma;
    if (t->h_block) cudaFreeHost(t->h_block);
    if (t->h_time) cudaFreeHost(t->h_time);
    if (t->h_wn) cudaFreeHost(t->h_wn);
    if (t->h_flag) cudaFreeHost(t->h_flag);
    if (t->h_lcur) cudaFreeHost(t->h_lcur);
    if (t->h_lnext) cudaFreeHost(t->h_lnext);
    if (t->h_wnext) cudaFreeHost(t->h_wnext);
    if (t->h_rstatus) cudaFreeHost(t->h_rstatus);
    for (cudaEvent_t e : t->prof.ev) cudaEventDestroy(e);
    if (t->ev0) cudaEventDestroy(t->ev0);
    if (t->ev1) cudaEventDestroy(t->ev1);
    if (t->stream) cudaStreamDestroy(t->stream);
    delete t;
}

}  // extern "C"

// =================================================================================================
// frame pipeline
// =================================================================================================
namespace vs {

int launch_ransac(vs_tracker* t);   // ransac.cu
// constant salt: the per-stream seed comes from the stream's points (ransac.cuh), so results do not depend on
// how streams are batched, split into lanes or spread over handles
static inline unsigned ransac_seed(const vs_tracker*) { return 12345u; }

static void fill_lk_params(vs_tracker* t, LkParams& p, int prev_buf, int next_buf, const float2* pts, float2* out,
                           uint8_t* status) {
    p.prev_pyr = t->d_pyr[prev_buf];
    p.next_pyr = t->d_pyr[next_buf];
    p.pd = t->pd;
    int nl = t->built[prev_buf] < t->built[next_buf] ? t->built[prev_buf] : t->built[next_buf];
    if (nl > t->cfg.lk_pyramid_levels + 1) nl = t->cfg.lk_pyramid_levels + 1;
    p.n_levels = nl;
    p.win = t->cfg.lk_window_size;
    int mc = t->cfg.lk_iterations;            // TermCriteria clamps inside calcOpticalFlowPyrLK
    mc = mc < 0 ? 0 : (mc > 100 ? 100 : mc);
    p.max_count = mc;
    double e = t->cfg.lk_eps;
    e = e < 0.0 ? 0.0 : (e > 10.0 ? 10.0 : e);
    p.eps2 = e * e;
    p.min_eig_threshold = 1e-4f;
    p.pts = pts;
    p.counts = t->g.n;
    p.cap = t->cap;
    p.next_pts = out;
    p.status = status;
    p.work_counters = t->d_lk_counters;
    p.work_seq = t->lk_seq++;
    p.tma_maps = t->lk_tma;
    p.prev_buf = prev_buf;
    p.next_buf = next_buf;
}

// build the pyramid of the new frame into buffer `buf` from a device image
static int build_pyramid(vs_tracker* t, int buf, const uint8_t* d_img, int pitch, int64_t stride) {
    const LevelDesc& l0 = t->pd.lv[0];
    if (t->cfg.equalize) {
        launch_clahe_lut_only(d_img, pitch, stride, t->W, t->H, t->B, t->d_luts, t->stream);
        prof_tick(t, PK_CLAHE_LUT);
        launch_clahe_apply_only(d_img, pitch, stride, t->d_pyr[buf] + l0.offset, l0.pitch, t->pd.stream_bytes, t->W,
                                t->H, t->B, t->d_luts, t->stream);
        prof_tick(t, PK_CLAHE_APPLY);
        t->launches += 2;
    } else {
        launch_copy_image(d_img, pitch, stride, t->d_pyr[buf] + l0.offset, l0.pitch, t->pd.stream_bytes, t->W, t->H,
                          t->B, t->stream, &t->launches);
        prof_tick(t, PK_CLAHE_APPLY);
    }
    const int nl = levels_to_build(t);
    for (int l = 0; l + 1 < nl; ++l) launch_pyr_down(t->d_pyr[buf], t->pd, l, t->B, t->stream, &t->launches);
    prof_tick(t, PK_PYR_DOWN);
    launch_pyr_pad(t->d_pyr[buf], t->pd, nl, t->B, t->stream, &t->launches);
    prof_tick(t, PK_PYR_PAD);
    t->built[buf] = nl;
    return VS_OK;
}

static int ransac_callback_hop(vs_tracker* t) {
    GlueParams& g = t->g;
    const size_t B = t->B, C = t->cap;
    VS_CUDA_TRY(cudaMemcpyAsync(t->h_wn, g.wn, sizeof(int) * B, cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(t->h_flag, g.ransac_flag, sizeof(int) * B, cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(t->h_lcur, g.w_lcur, sizeof(float2) * B * C, cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(t->h_lnext, g.w_lnext, sizeof(float2) * B * C, cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(t->h_wnext, g.w_next, sizeof(float2) * B * C, cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
    bool any = false;
    for (size_t s = 0; s < B; ++s) {
        if (!t->h_flag[s]) continue;
        any = true;
        const int n = t->h_wn[s];
        uint8_t* st = t->h_rstatus + s * C;
        memset(st, 1, n);
        if (t->reject_fn) {
            const int rc = t->reject_fn(t->reject_user, (int)s, n, reinterpret_cast<const float*>(t->h_lcur + s * C),
                                        reinterpret_cast<const float*>(t->h_lnext + s * C),
                                        reinterpret_cast<const float*>(t->h_wnext + s * C), st);
            if (rc != 0) { set_error("reject callback failed for stream %d (rc=%d)", (int)s, rc); return VS_ERR_STATE; }
        }
    }
    if (any) VS_CUDA_TRY(cudaMemcpyAsync(g.w_rstatus, t->h_rstatus, B * C, cudaMemcpyHostToDevice, t->stream));
    return VS_OK;
}

// Shi-Tomasi launch parameters: image = level 0 of pyramid buffer `buf`, mask discs = g's kept set
static StParams make_st_params(vs_tracker* t, int buf, const GlueParams& g) {
    StParams e{};
    e.img = t->d_pyr[buf] + t->pd.lv[0].offset;
    e.pitch = t->pd.lv[0].pitch;
    e.img_stride = t->pd.stream_bytes;
    e.W = t->W;
    e.H = t->H;
    e.want = g.want;
    e.kn = g.kn;
    e.k_pts = g.k_pts;
    e.cap = g.cap;
    e.min_dist = g.min_dist;
    e.disc_hw = g.disc_hw;
    e.base_mode = t->base_mode;
    e.fcx = t->fcx;
    e.fcy = t->fcy;
    e.fr = t->fr;
    e.fish_hw = t->d_fish_hw;
    e.base_mask = g.base_mask;
    e.mask_pitch = g.mask_pitch;
    e.max_enc = g.max_enc;
    e.cand = g.cand;
    e.cand_count = g.cand_count;
    e.cand_cap = g.cand_cap;
    e.raw = g.cand_raw;
    e.raw_count = g.raw_count;
    e.l_enc = g.l_enc;
    e.eig_out = nullptr;
    e.mask_out = nullptr;
    e.quality = 0.01;
    return e;
}

static int run_frame(vs_tracker* t, const uint8_t* d_img, int pitch, int64_t stride) {
    GlueParams& g = t->g;
    sync_glue_params(t);
    const int rc0 = upload_disc(t);
    if (rc0 != VS_OK) return rc0;
    if (t->n_bound < t->cfg.max_cnt) t->n_bound = t->cfg.max_cnt;
    const int nb = 1 - t->cur;
    build_pyramid(t, nb, d_img, pitch, stride);
    if (t->has_prev) {
        LkParams p;
        fill_lk_params(t, p, t->cur, nb, g.pts, g.lk_next, g.lk_status);
        if (launch_lk(p, t->B, t->n_bound, t->stream, &t->launches) != 0) { set_error("LK: unsupported window"); return VS_ERR_INVALID; }
        prof_tick(t, PK_LK);
        if (t->cfg.fb_check) {
            LkParams q;
            fill_lk_params(t, q, nb, t->cur, g.lk_next, g.lk_back, g.lk_back_status);
            launch_lk(q, t->B, t->n_bound, t->stream, &t->launches);
            prof_tick(t, PK_LK_BACK);
        }
    }
    // post_lk (+ device RANSAC) + set_mask in one launch; the profiled pass keeps RANSAC as its own launch so that
    // its share of the frame stays visible
    const bool fuse_ransac = t->cfg.ransac_mode == VS_RANSAC_DEVICE && !t->prof.on;
    const bool fuse_glue = t->cfg.ransac_mode == VS_RANSAC_OFF || fuse_ransac;
    if (!fuse_glue) {
        launch_post_lk(g, t->stream, &t->launches);
        prof_tick(t, PK_POST_LK);
    }
    if (t->cfg.ransac_mode == VS_RANSAC_CALLBACK) {
        const int rc = ransac_callback_hop(t);
        if (rc != VS_OK) return rc;
    } else if (t->cfg.ransac_mode == VS_RANSAC_DEVICE && !fuse_ransac) {
        const int rc = launch_ransac(t);
        if (rc != VS_OK) return rc;
    }
    if (!fuse_glue) prof_tick(t, PK_RANSAC);
    launch_set_mask(g, t->stream, &t->launches, fuse_glue, fuse_ransac, ransac_seed(t));
    prof_tick(t, PK_SET_MASK);
    StParams e = make_st_params(t, nb, g);
    e.img_padded = kPyrFirstPaddedLevel == 0 && kPyrPadX >= 8;       // build_pyramid wrote level 0's border columns
    launch_shi_tomasi(e, t->B, t->stream, &t->launches);
    prof_tick(t, PK_SHI_TOMASI);
    launch_select_finish(g, t->stream, &t->launches);
    prof_tick(t, PK_SELECT_FINISH);
    t->cur = nb;
    t->has_prev = true;
    VS_CUDA_TRY(cudaGetLastError());
    return VS_OK;
}

static void fill_results(vs_tracker* t, vs_track_result* out, int slot = 0) {
    const int64_t B = t->B, C = t->cap;
    const uint8_t* p = t->h_block + (size_t)slot * t->h_block_bytes;
    const int* hdr = reinterpret_cast<const int*>(p); p += B * kHdrInts * 4;
    const int* ids = reinterpret_cast<const int*>(p); p += B * C * 4;
    const int* cnt = reinterpret_cast<const int*>(p); p += B * C * 4;
    const float* pts = reinterpret_cast<const float*>(p); p += B * C * 8;
    const float* un = reinterpret_cast<const float*>(p); p += B * C * 8;
    const float* vel = reinterpret_cast<const float*>(p);
    for (int64_t s = 0; s < B; ++s) {
        const int* h = hdr + s * kHdrInts;
        vs_track_result& r = out[s];
        r.n = h[H_N];
        r.n_new = h[H_NNEW];
        r.ids = ids + s * C;
        r.track_cnt = cnt + s * C;
        r.cur_pts = pts + s * C * 2;
        r.undist = un + s * C * 2;
        r.velocity = vel + s * C * 2;
        r.n_lk_in = h[H_LK_IN];
        r.n_lk_ok = h[H_LK_OK];
        r.n_ransac_in = h[H_RANSAC_IN];
        r.n_ransac_out = h[H_RANSAC_OUT];
        r.n_mask_kept = h[H_MASK_KEPT];
        r.n_candidates = h[H_NCAND];
        r.ransac_ran = h[H_RANSAC_RAN];
        float f;
        memcpy(&f, &h[H_RMS_BITS], 4);
        r.rms_disp = f;
    }
}

}  // namespace vs

namespace vs {

// ---- asynchronous halves of a frame call (everything is enqueued on t->stream; nothing waits) ----

// host frames -> d_raw.  Tightly packed frames go up as 1-D copies, and runs of frames that are also
// consecutive in host memory as ONE copy (a 2-D copy per frame costs a driver call each and, for widths
// that are not a multiple of 64, an order of magnitude in PCIe efficiency).
static int enqueue_upload(vs_tracker* t, const uint8_t* const* frames, int stride) {
    const int64_t img_bytes = (int64_t)t->W * t->H;
    const bool packed = stride == t->W && t->raw_pitch == t->W && t->raw_stride == img_bytes;
    for (int s = 0; s < t->B; ++s)
        if (!frames[s]) { set_error("frames[%d] is null", s); return VS_ERR_INVALID; }
    if (packed) {
        int s = 0;
        while (s < t->B) {
            int e = s + 1;
            while (e < t->B && frames[e] == frames[e - 1] + img_bytes) ++e;
            VS_CUDA_TRY(cudaMemcpyAsync(t->d_raw + (int64_t)s * t->raw_stride, frames[s], (size_t)(img_bytes * (e - s)),
                                        cudaMemcpyHostToDevice, t->stream));
            s = e;
        }
    } else {
        for (int s = 0; s < t->B; ++s)
            VS_CUDA_TRY(cudaMemcpy2DAsync(t->d_raw + (int64_t)s * t->raw_stride, t->raw_pitch, frames[s], stride, t->W,
                                          t->H, cudaMemcpyHostToDevice, t->stream));
    }
    return VS_OK;
}

// one frame per stream, images resident on the device
static int enqueue_frame(vs_tracker* t, const uint8_t* d_frames, int32_t stride, int64_t image_stride_bytes,
                         const double* time_s, int slot = 0) {
    // stream-ordered timestamps: slot k of the ring is rewritten only after the frame that read it has finished
    // (calls with results_to_host = 0 return without a sync, so a fixed slot would be overwritten under a running frame)
    const int ts = (int)(t->time_seq % vs_tracker::kTimeRing);
    if (t->time_seq >= vs_tracker::kTimeRing) VS_CUDA_TRY(cudaEventSynchronize(t->ev_time[ts]));
    double* h_time = t->h_time + (size_t)ts * t->B;
    memcpy(h_time, time_s, sizeof(double) * t->B);
    static const int pdl_max = [] { const char* e = getenv("VS_PDL_MAX_STREAMS"); return e ? atoi(e) : kPdlMaxStreams; }();
    tl_pdl_launch = t->B <= pdl_max;              // programmatic dependent launches for latency-sized handles only
    VS_CUDA_TRY(cudaEventRecord(t->ev0, t->stream));
    prof_begin(t);
    t->g.cur_time = h_time;      // pinned host memory is device-addressable (UVA): select_finish reads the B doubles in place
    const uint8_t* img = d_frames;
    int pitch = stride;
    int64_t istride = image_stride_bytes;
    if ((stride & 15) || (reinterpret_cast<uintptr_t>(d_frames) & 15) || (image_stride_bytes & 15)) {
        // kernels use aligned vector loads: stage unaligned device input through the internal buffer
        launch_copy_image(d_frames, stride, image_stride_bytes, t->d_raw, t->raw_pitch, t->raw_stride, t->W, t->H, t->B,
                          t->stream, &t->launches);
        img = t->d_raw;
        pitch = t->raw_pitch;
        istride = t->raw_stride;
    }
    prof_tick(t, PK_INPUT);
    if (direct_results(t)) point_results(t, t->h_block + (size_t)slot * t->h_block_bytes);
    t->last_results = direct_results(t) ? t->h_block + (size_t)slot * t->h_block_bytes : t->d_rblock;
    const int rc = run_frame(t, img, pitch, istride);
    if (direct_results(t)) point_results(t, t->d_rblock);      // stage entry points and fetches use the device block
    if (rc != VS_OK) return rc;
    VS_CUDA_TRY(cudaEventRecord(t->ev1, t->stream));
    VS_CUDA_TRY(cudaEventRecord(t->ev_time[ts], t->stream));
    ++t->time_seq;
    prof_end(t);
    return VS_OK;
}

static int enqueue_fetch(vs_tracker* t, int slot = 0) {
    if (direct_results(t)) {            // the frame's kernels already wrote the slot
        VS_CUDA_TRY(cudaEventRecord(t->ev_fetch[slot], t->stream));
        return VS_OK;
    }
    VS_CUDA_TRY(cudaMemcpyAsync(t->h_block + (size_t)slot * t->h_block_bytes, t->d_rblock, t->h_block_bytes,
                                cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaEventRecord(t->ev_fetch[slot], t->stream));
    return VS_OK;
}

// lanes: make every lane's stream wait for work already queued on the parent stream (fork) ...
static int lanes_fork(vs_tracker* t) {
    VS_CUDA_TRY(cudaEventRecord(t->ev0, t->stream));
    for (vs_tracker* lane : t->lanes) VS_CUDA_TRY(cudaStreamWaitEvent(lane->stream, t->ev0, 0));
    return VS_OK;
}
// ... and the parent stream wait for every lane (join), so events recorded on the parent stream bracket the call
static int lanes_join(vs_tracker* t) {
    for (vs_tracker* lane : t->lanes) {
        VS_CUDA_TRY(cudaEventRecord(lane->ev_done, lane->stream));
        VS_CUDA_TRY(cudaStreamWaitEvent(t->stream, lane->ev_done, 0));
    }
    VS_CUDA_TRY(cudaEventRecord(t->ev1, t->stream));
    return VS_OK;
}

}  // namespace vs

extern "C" {

int vs_tracker_fetch_results(vs_tracker* t, vs_track_result* out) {
    if (!t || !out) { set_error("null argument"); return VS_ERR_INVALID; }
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    if (!t->lanes.empty()) {
        for (vs_tracker* lane : t->lanes) { const int rc = enqueue_fetch(lane); if (rc != VS_OK) return rc; }
        for (size_t l = 0; l < t->lanes.size(); ++l) {
            VS_CUDA_TRY(cudaStreamSynchronize(t->lanes[l]->stream));
            fill_results(t->lanes[l], out + t->lane_s0[l]);
        }
        return VS_OK;
    }
    const int rc = enqueue_fetch(t);
    if (rc != VS_OK) return rc;
    VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
    fill_results(t, out);
    return VS_OK;
}

int vs_tracker_track_batch_device(vs_tracker* t, const uint8_t* d_frames, int32_t stride, int64_t image_stride_bytes,
                                  const double* time_s, int32_t results_to_host, vs_track_result* out) {
    if (!t || !d_frames || !time_s) { set_error("null argument"); return VS_ERR_INVALID; }
    if (t->submit_seq != t->wait_seq) { set_error("a submitted batch is still in flight: call vs_tracker_wait first"); return VS_ERR_STATE; }
    if (stride < t->W) { set_error("stride < width"); return VS_ERR_INVALID; }
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    if (!t->lanes.empty()) {
        int rc = lanes_fork(t);
        if (rc != VS_OK) return rc;
        for (size_t l = 0; l < t->lanes.size(); ++l) {
            const int s0 = t->lane_s0[l];
            rc = enqueue_frame(t->lanes[l], d_frames + (int64_t)s0 * image_stride_bytes, stride, image_stride_bytes, time_s + s0);
            if (rc != VS_OK) return rc;
        }
        rc = lanes_join(t);
        if (rc != VS_OK) return rc;
        if (results_to_host && out) {
            rc = vs_tracker_fetch_results(t, out);
            if (rc != VS_OK) return rc;
            VS_CUDA_TRY(cudaEventSynchronize(t->ev1));
            cudaEventElapsedTime(&t->last_ms, t->ev0, t->ev1);
        }
        return VS_OK;
    }
    const int rc = enqueue_frame(t, d_frames, stride, image_stride_bytes, time_s);
    if (rc != VS_OK) return rc;
    if (results_to_host && out) {
        const int rc2 = vs_tracker_fetch_results(t, out);
        if (rc2 != VS_OK) return rc2;
        cudaEventElapsedTime(&t->last_ms, t->ev0, t->ev1);
    }
    return VS_OK;
}

int vs_tracker_track_batch(vs_tracker* t, const uint8_t* const* frames, int32_t stride, const double* time_s,
                           vs_track_result* out) {
    if (!t || !frames || !time_s || !out) { set_error("null argument"); return VS_ERR_INVALID; }
    if (t->submit_seq != t->wait_seq) { set_error("a submitted batch is still in flight: call vs_tracker_wait first"); return VS_ERR_STATE; }
    if (stride < t->W) { set_error("stride < width"); return VS_ERR_INVALID; }
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    if (!t->lanes.empty()) {
        // lane l: upload -> kernels -> result download on its own stream; the copy engines work on lane l+1
        // while the SMs run lane l
        int rc = lanes_fork(t);
        if (rc != VS_OK) return rc;
        for (size_t l = 0; l < t->lanes.size(); ++l) {
            vs_tracker* lane = t->lanes[l];
            const int s0 = t->lane_s0[l];
            rc = enqueue_upload(lane, frames + s0, stride);
            if (rc == VS_OK) rc = enqueue_frame(lane, lane->d_raw, lane->raw_pitch, lane->raw_stride, time_s + s0);
            if (rc == VS_OK) rc = enqueue_fetch(lane);
            if (rc != VS_OK) return rc;
        }
        rc = lanes_join(t);
        if (rc != VS_OK) return rc;
        for (size_t l = 0; l < t->lanes.size(); ++l) {
            VS_CUDA_TRY(cudaStreamSynchronize(t->lanes[l]->stream));
            fill_results(t->lanes[l], out + t->lane_s0[l]);
        }
        VS_CUDA_TRY(cudaEventSynchronize(t->ev1));
        cudaEventElapsedTime(&t->last_ms, t->ev0, t->ev1);
        return VS_OK;
    }
    int rc = enqueue_upload(t, frames, stride);
    if (rc != VS_OK) return rc;
    return vs_tracker_track_batch_device(t, t->d_raw, t->raw_pitch, t->raw_stride, time_s, 1, out);
}

// ---- asynchronous pair: up to two batches in flight, so the upload of batch k+1 overlaps the kernels of batch k ----
int vs_tracker_submit_batch(vs_tracker* t, const uint8_t* const* frames, int32_t stride, const double* time_s) {
    if (!t || !frames || !time_s) { set_error("null argument"); return VS_ERR_INVALID; }
    if (stride < t->W) { set_error("stride < width"); return VS_ERR_INVALID; }
    if (t->submit_seq - t->wait_seq >= 2) { set_error("two batches already in flight: call vs_tracker_wait first"); return VS_ERR_STATE; }
    if (t->cfg.ransac_mode == VS_RANSAC_CALLBACK) { set_error("submit/wait is not available with VS_RANSAC_CALLBACK"); return VS_ERR_STATE; }
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    const int slot = (int)(t->submit_seq & 1);
    if (!t->lanes.empty()) {
        for (size_t l = 0; l < t->lanes.size(); ++l) {
            vs_tracker* lane = t->lanes[l];
            const int s0 = t->lane_s0[l];
            int rc = enqueue_upload(lane, frames + s0, stride);
            if (rc == VS_OK) rc = enqueue_frame(lane, lane->d_raw, lane->raw_pitch, lane->raw_stride, time_s + s0, slot);
            if (rc == VS_OK) rc = enqueue_fetch(lane, slot);
            if (rc != VS_OK) return rc;
        }
    } else {
        int rc = enqueue_upload(t, frames, stride);
        if (rc == VS_OK) rc = enqueue_frame(t, t->d_raw, t->raw_pitch, t->raw_stride, time_s, slot);
        if (rc == VS_OK) rc = enqueue_fetch(t, slot);
        if (rc != VS_OK) return rc;
    }
    ++t->submit_seq;
    return VS_OK;
}

int vs_tracker_wait(vs_tracker* t, vs_track_result* out) {
    if (!t || !out) { set_error("null argument"); return VS_ERR_INVALID; }
    if (t->submit_seq == t->wait_seq) { set_error("vs_tracker_wait: nothing in flight"); return VS_ERR_STATE; }
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    const int slot = (int)(t->wait_seq & 1);
    if (!t->lanes.empty()) {
        for (size_t l = 0; l < t->lanes.size(); ++l) {
            VS_CUDA_TRY(cudaEventSynchronize(t->lanes[l]->ev_fetch[slot]));
            fill_results(t->lanes[l], out + t->lane_s0[l], slot);
        }
    } else {
        VS_CUDA_TRY(cudaEventSynchronize(t->ev_fetch[slot]));
        fill_results(t, out, slot);
    }
    ++t->wait_seq;
    return VS_OK;
}

int vs_tracker_track(vs_tracker* t, const uint8_t* gray, int32_t stride, double time_s, vs_track_result* out) {
    if (!t) { set_error("null tracker"); return VS_ERR_INVALID; }
    if (t->B != 1) { set_error("vs_tracker_track needs n_streams == 1 (use vs_tracker_track_batch)"); return VS_ERR_STATE; }
    const uint8_t* frames[1] = {gray};
    return vs_tracker_track_batch(t, frames, stride, &time_s, out);
}

int vs_tracker_set_params(vs_tracker* t, int32_t max_cnt, int32_t min_dist, double f_threshold, int32_t lk_window,
                          int32_t lk_pyramid, int32_t lk_iterations, double lk_eps, double f_edge_factor) {
    if (!t) { set_error("null tracker"); return VS_ERR_INVALID; }
    for (vs_tracker* lane : t->lanes) {
        const int rc = vs_tracker_set_params(lane, max_cnt, min_dist, f_threshold, lk_window, lk_pyramid, lk_iterations,
                                             lk_eps, f_edge_factor);
        if (rc != VS_OK) return rc;
    }
    if (max_cnt > 0) {
        if (max_cnt > t->cap) { set_error("max_cnt %d above handle capacity %d", max_cnt, t->cap); return VS_ERR_CAPACITY; }
        t->cfg.max_cnt = max_cnt;
    }
    if (min_dist > 0) {
        if (min_dist > kMaxDiscRadius) { set_error("min_dist too large"); return VS_ERR_INVALID; }
        t->cfg.min_dist = min_dist;
    }
    if (f_threshold > 0.0) t->cfg.f_threshold = f_threshold;
    if (lk_window > 0 && (lk_window % 2 == 1)) {       // same guard as VIOEngine::setTrackingParams
        if (lk_window > kMaxWin || lk_window < 3) { set_error("lk_window above %d unsupported", kMaxWin); return VS_ERR_INVALID; }
        t->cfg.lk_window_size = lk_window;
    }
    if (lk_pyramid > 0 && lk_pyramid <= 5) t->cfg.lk_pyramid_levels = lk_pyramid;
    if (lk_iterations > 0) t->cfg.lk_iterations = lk_iterations;
    if (lk_eps > 0.0) t->cfg.lk_eps = lk_eps;
    if (f_edge_factor >= 0.0) t->cfg.f_threshold_edge_factor = f_edge_factor;
    return VS_OK;
}

int vs_tracker_set_reject_callback(vs_tracker* t, vs_reject_fn fn, void* user) {
    if (!t) { set_error("null tracker"); return VS_ERR_INVALID; }
    t->reject_fn = fn;
    t->reject_user = user;
    return VS_OK;
}

int vs_tracker_set_frontend_flags(vs_tracker* t, int32_t equalize, int32_t fisheye, double focal_length,
                                  const uint8_t* fisheye_mask_image, int32_t fisheye_mask_stride) {
    if (!t) { set_error("null tracker"); return VS_ERR_INVALID; }
    if (t->submit_seq != t->wait_seq) { set_error("a submitted batch is still in flight: call vs_tracker_wait first"); return VS_ERR_STATE; }
    for (vs_tracker* lane : t->lanes) {
        const int rc = vs_tracker_set_frontend_flags(lane, equalize, fisheye, focal_length, fisheye_mask_image, fisheye_mask_stride);
        if (rc != VS_OK) return rc;
    }
    if (equalize >= 0) t->cfg.equalize = equalize ? 1 : 0;          // read by the next frame's build_pyramid
    if (focal_length > 0.0) t->cfg.focal_length = focal_length;       // read by the next frame's sync_glue_params
    if (t->lanes.empty() && (fisheye >= 0 || fisheye_mask_image)) {
        const int want = fisheye >= 0 ? (fisheye ? 1 : 0) : t->cfg.fisheye;
        if (want != t->cfg.fisheye || fisheye_mask_image) {
            VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
            VS_CUDA_TRY(cudaStreamSynchronize(t->stream));             // frames in flight still read the old mask
            const int rc = build_base_mask(t, want, fisheye_mask_image, fisheye_mask_stride);
            if (rc != VS_OK) return rc;
        }
    } else if (fisheye >= 0) {
        t->cfg.fisheye = fisheye ? 1 : 0;
    }
    return VS_OK;
}

int vs_tracker_get_config(const vs_tracker* t, vs_tracker_config* out) {
    if (!t || !out) { set_error("null argument"); return VS_ERR_INVALID; }
    *out = t->cfg;
    if (!t->lanes.empty()) {                       // the parent of a laned handle forwards set_params to its lanes
        const vs_tracker_config& l = t->lanes[0]->cfg;
        out->max_cnt = l.max_cnt; out->min_dist = l.min_dist; out->f_threshold = l.f_threshold;
        out->lk_window_size = l.lk_window_size; out->lk_pyramid_levels = l.lk_pyramid_levels;
        out->lk_iterations = l.lk_iterations; out->lk_eps = l.lk_eps; out->f_threshold_edge_factor = l.f_threshold_edge_factor;
    }
    out->fisheye_mask_image = nullptr;
    return VS_OK;
}

int vs_tracker_reset(vs_tracker* t, int32_t stream) {
    if (!t) { set_error("null tracker"); return VS_ERR_INVALID; }
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    if (!t->lanes.empty()) {
        if (stream >= t->B) { set_error("bad stream"); return VS_ERR_INVALID; }
        for (size_t l = 0; l < t->lanes.size(); ++l) {
            const int s0 = t->lane_s0[l], s1 = t->lane_s0[l + 1];
            if (stream < 0 || (stream >= s0 && stream < s1)) {
                const int rc = vs_tracker_reset(t->lanes[l], stream < 0 ? -1 : stream - s0);
                if (rc != VS_OK) return rc;
            }
        }
        return VS_OK;
    }
    if (stream < 0) {
        VS_CUDA_TRY(cudaMemsetAsync(t->g.n, 0, sizeof(int) * t->B, t->stream));
        t->has_prev = false;
    } else {
        if (stream >= t->B) { set_error("bad stream"); return VS_ERR_INVALID; }
        VS_CUDA_TRY(cudaMemsetAsync(t->g.n + stream, 0, sizeof(int), t->stream));
    }
    VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
    return VS_OK;
}

int64_t vs_tracker_debug_guard_check(vs_tracker* t) {
    if (!t) return -2;
    if (cudaSetDevice(t->cfg.device) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) return -2;
    if (t->lanes.empty()) return t->arena.check();
    int64_t bad = 0;
    for (vs_tracker* l : t->lanes) {
        const int64_t b = l->arena.check();
        if (b < 0) return b;
        bad += b;
    }
    return bad;
}

int vs_tracker_debug_read_pyramid(vs_tracker* t, int32_t stream, uint8_t* out, int64_t cap_bytes, int64_t* layout, int32_t* n_levels) {
    if (!t || !layout || !n_levels) { set_error("vs_tracker_debug_read_pyramid: null argument"); return VS_ERR_INVALID; }
    if (!t->lanes.empty()) {
        int l = 0;
        while (l + 1 < (int)t->lanes.size() && stream >= t->lane_s0[l + 1]) ++l;
        stream -= t->lane_s0[l];
        t = t->lanes[l];
    }
    if (stream < 0 || stream >= t->B) { set_error("vs_tracker_debug_read_pyramid: bad stream"); return VS_ERR_INVALID; }
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    const int nl = t->built[t->cur];
    for (int l = 0; l < kMaxLevels; ++l) {
        layout[4 * l + 0] = t->pd.lv[l].offset;
        layout[4 * l + 1] = t->pd.lv[l].pitch;
        layout[4 * l + 2] = t->pd.lv[l].w;
        layout[4 * l + 3] = t->pd.lv[l].h;
    }
    layout[4 * kMaxLevels] = t->pd.stream_bytes;
    layout[4 * kMaxLevels + 1] = kPyrPadX;
    layout[4 * kMaxLevels + 2] = kPyrPadY;
    *n_levels = nl;
    if (out) {
        if (cap_bytes < t->pd.stream_bytes) { set_error("vs_tracker_debug_read_pyramid: buffer below %lld bytes", (long long)t->pd.stream_bytes); return VS_ERR_CAPACITY; }
        VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
        VS_CUDA_TRY(cudaMemcpy(out, t->d_pyr[t->cur] + (int64_t)stream * t->pd.stream_bytes, (size_t)t->pd.stream_bytes, cudaMemcpyDeviceToHost));
    }
    return VS_OK;
}

int64_t vs_tracker_launch_count(const vs_tracker* t) {
    if (!t) return 0;
    int64_t n = t->launches;
    for (const vs_tracker* lane : t->lanes) n += lane->launches;
    return n;
}
void* vs_tracker_cuda_stream(const vs_tracker* t) { return t ? (void*)t->stream : nullptr; }

int vs_tracker_set_profiling(vs_tracker* t, int32_t on) {
    if (!t) { set_error("null tracker"); return VS_ERR_INVALID; }
    for (vs_tracker* lane : t->lanes) vs_tracker_set_profiling(lane, on);
    t->prof.on = on != 0;
    if (on)
        for (int i = 0; i < PK_COUNT; ++i) { t->prof.total_ms[i] = 0.0; t->prof.launches[i] = 0; }
    return VS_OK;
}

int vs_tracker_get_profile(const vs_tracker* t, int32_t max_n, float* total_ms, int32_t* launches, int32_t* n_out) {
    if (!t || !total_ms || !launches || !n_out) { set_error("null argument"); return VS_ERR_INVALID; }
    const int n = max_n < PK_COUNT ? max_n : (int)PK_COUNT;
    for (int i = 0; i < n; ++i) { total_ms[i] = (float)t->prof.total_ms[i]; launches[i] = t->prof.launches[i]; }
    for (const vs_tracker* lane : t->lanes)
        for (int i = 0; i < n; ++i) { total_ms[i] += (float)lane->prof.total_ms[i]; launches[i] += lane->prof.launches[i]; }
    *n_out = n;
    return VS_OK;
}

int vs_pack_image_features(const vs_track_result* r, int32_t max_n, int32_t* ids_out, double* rec_out, int32_t* n_out) {
    if (!r || !ids_out || !rec_out || !n_out || max_n < 0) { set_error("vs_pack_image_features: bad argument"); return VS_ERR_INVALID; }
    int n = 0, rc = VS_OK;
    for (int j = 0; j < r->n; ++j) {
        if (r->track_cnt[j] <= 1) continue;                     // measurement_processor.cpp:211
        if (n >= max_n) { set_error("vs_pack_image_features: more than %d published features", max_n); rc = VS_ERR_CAPACITY; break; }
        ids_out[n] = r->ids[j];
        double* o = rec_out + 7 * (size_t)n;
        o[0] = r->undist[2 * j];   o[1] = r->undist[2 * j + 1];   o[2] = 1.0;
        o[3] = r->cur_pts[2 * j];  o[4] = r->cur_pts[2 * j + 1];
        o[5] = r->velocity[2 * j]; o[6] = r->velocity[2 * j + 1];
        ++n;
    }
    *n_out = n;
    return rc;
}

const char* vs_profile_kernel_name(int32_t slot) { return (slot >= 0 && slot < PK_COUNT) ? kProfNames[slot] : ""; }
float vs_tracker_last_device_ms(const vs_tracker* t) { return t ? t->last_ms : 0.f; }

}  // extern "C"

// =================================================================================================
// stage entry points — the same kernels as the frame path, on stream slot 0 of the handle.
// They use the handle's buffers as scratch, so tracking state is reset afterwards.
// =================================================================================================
namespace vs {

static int stage_begin(vs_tracker*& t) {
    if (!t) { set_error("null tracker"); return VS_ERR_INVALID; }
    if (!t->lanes.empty()) t = t->lanes[0];      // stage entry points use one stream slot: lane 0 is a full handle
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    sync_glue_params(t);
    return upload_disc(t);
}

static int stage_end(vs_tracker* t) {
    VS_CUDA_TRY(cudaMemsetAsync(t->g.n, 0, sizeof(int) * t->B, t->stream));
    VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
    VS_CUDA_TRY(cudaGetLastError());
    t->has_prev = false;
    return VS_OK;
}

static int upload_level0(vs_tracker* t, int buf, const uint8_t* img) {
    const LevelDesc& l0 = t->pd.lv[0];
    VS_CUDA_TRY(cudaMemcpy2DAsync(t->d_pyr[buf] + l0.offset, l0.pitch, img, t->W, t->W, t->H, cudaMemcpyHostToDevice,
                                  t->stream));
    return VS_OK;
}

}  // namespace vs

extern "C" {

int vs_stage_clahe(vs_tracker* t, const uint8_t* img, uint8_t* out) {
    int rc = stage_begin(t);
    if (rc != VS_OK) return rc;
    if (!img || !out) { set_error("null argument"); return VS_ERR_INVALID; }
    const int nb = 1 - t->cur;
    const LevelDesc& l0 = t->pd.lv[0];
    VS_CUDA_TRY(cudaMemcpy2DAsync(t->d_raw, t->raw_pitch, img, t->W, t->W, t->H, cudaMemcpyHostToDevice, t->stream));
    launch_clahe(t->d_raw, t->raw_pitch, t->raw_stride, t->d_pyr[nb] + l0.offset, l0.pitch, t->pd.stream_bytes, t->W,
                 t->H, 1, t->d_luts, t->stream, &t->launches);
    VS_CUDA_TRY(cudaMemcpy2DAsync(out, t->W, t->d_pyr[nb] + l0.offset, l0.pitch, t->W, t->H, cudaMemcpyDeviceToHost,
                                  t->stream));
    return stage_end(t);
}

int vs_stage_pyramid(vs_tracker* t, const uint8_t* level0, uint8_t* out_levels, int32_t* sizes, int32_t* n_levels) {
    int rc = stage_begin(t);
    if (rc != VS_OK) return rc;
    if (!level0 || !out_levels || !sizes || !n_levels) { set_error("null argument"); return VS_ERR_INVALID; }
    const int nb = 1 - t->cur;
    rc = upload_level0(t, nb, level0);
    if (rc != VS_OK) return rc;
    const int nl = levels_to_build(t);
    for (int l = 0; l + 1 < nl; ++l) launch_pyr_down(t->d_pyr[nb], t->pd, l, 1, t->stream, &t->launches);
    size_t off = 0;
    for (int l = 0; l < nl; ++l) {
        const LevelDesc& L = t->pd.lv[l];
        VS_CUDA_TRY(cudaMemcpy2DAsync(out_levels + off, L.w, t->d_pyr[nb] + L.offset, L.pitch, L.w, L.h,
                                      cudaMemcpyDeviceToHost, t->stream));
        off += (size_t)L.w * L.h;
        sizes[2 * l] = L.w;
        sizes[2 * l + 1] = L.h;
    }
    *n_levels = nl;
    return stage_end(t);
}

int vs_stage_lk(vs_tracker* t, const uint8_t* prev_level0, const uint8_t* next_level0, const float* pts_xy, int32_t n,
                float* next_xy, uint8_t* status) {
    int rc = stage_begin(t);
    if (rc != VS_OK) return rc;
    if (!prev_level0 || !next_level0 || !pts_xy || !next_xy || !status || n < 0) { set_error("bad argument"); return VS_ERR_INVALID; }
    if (n > t->cap) { set_error("n %d above capacity %d", n, t->cap); return VS_ERR_CAPACITY; }
    const int nl = levels_to_build(t);
    for (int b = 0; b < 2; ++b) {
        rc = upload_level0(t, b, b == 0 ? prev_level0 : next_level0);
        if (rc != VS_OK) return rc;
        for (int l = 0; l + 1 < nl; ++l) launch_pyr_down(t->d_pyr[b], t->pd, l, 1, t->stream, &t->launches);
        launch_pyr_pad(t->d_pyr[b], t->pd, nl, 1, t->stream, &t->launches);
        t->built[b] = nl;
    }
    VS_CUDA_TRY(cudaMemcpyAsync(t->g.pts, pts_xy, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(t->g.n, &n, sizeof(int), cudaMemcpyHostToDevice, t->stream));
    LkParams p;
    fill_lk_params(t, p, 0, 1, t->g.pts, t->g.lk_next, t->g.lk_status);
    if (launch_lk(p, 1, n, t->stream, &t->launches) != 0) { set_error("LK: unsupported window"); return VS_ERR_INVALID; }
    VS_CUDA_TRY(cudaMemcpyAsync(next_xy, t->g.lk_next, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(status, t->g.lk_status, n, cudaMemcpyDeviceToHost, t->stream));
    return stage_end(t);
}

static int stage_prepare_mask(vs_tracker* t, const float* kept_xy, int n_kept, int want) {
    GlueParams g1 = t->g;
    g1.n_streams = 1;
    if (n_kept > 0) VS_CUDA_TRY(cudaMemcpyAsync(g1.k_pts, kept_xy, sizeof(float) * 2 * n_kept, cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(g1.kn, &n_kept, sizeof(int), cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(g1.want, &want, sizeof(int), cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(g1.cand_count, 0, sizeof(int), t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(g1.max_enc, 0, sizeof(int), t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(g1.raw_count, 0, sizeof(int), t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(g1.l_enc, 0, sizeof(int), t->stream));
    return VS_OK;
}

int vs_stage_min_eig(vs_tracker* t, const uint8_t* img, float* eig) {
    int rc = stage_begin(t);
    if (rc != VS_OK) return rc;
    if (!img || !eig) { set_error("null argument"); return VS_ERR_INVALID; }
    const int nb = 1 - t->cur;
    rc = upload_level0(t, nb, img);
    if (rc != VS_OK) return rc;
    if (!t->d_eig) VS_CUDA_TRY(cudaMalloc(&t->d_eig, sizeof(float) * (size_t)t->W * t->H));
    rc = stage_prepare_mask(t, nullptr, 0, 1);
    if (rc != VS_OK) return rc;
    StParams e = make_st_params(t, nb, t->g);
    e.eig_out = t->d_eig;
    launch_shi_tomasi(e, 1, t->stream, &t->launches);
    VS_CUDA_TRY(cudaMemcpyAsync(eig, t->d_eig, sizeof(float) * (size_t)t->W * t->H, cudaMemcpyDeviceToHost, t->stream));
    return stage_end(t);
}

int vs_stage_good_features(vs_tracker* t, const uint8_t* img, const float* kept_xy, int32_t n_kept, int32_t max_corners,
                           float* corners_xy, int32_t* n_out, uint8_t* mask_out) {
    int rc = stage_begin(t);
    if (rc != VS_OK) return rc;
    if (!img || !corners_xy || !n_out || n_kept < 0 || (n_kept > 0 && !kept_xy)) { set_error("bad argument"); return VS_ERR_INVALID; }
    if (n_kept + max_corners > t->cap) { set_error("n_kept + max_corners above capacity %d", t->cap); return VS_ERR_CAPACITY; }
    const int nb = 1 - t->cur;
    rc = upload_level0(t, nb, img);
    if (rc != VS_OK) return rc;
    rc = stage_prepare_mask(t, kept_xy, n_kept, max_corners);
    if (rc != VS_OK) return rc;
    StParams e = make_st_params(t, nb, t->g);
    if (mask_out) {
        e.mask_out = t->g.mask;
        e.mask_out_pitch = t->g.mask_pitch;
        e.mask_out_stride = t->g.mask_stride;
        VS_CUDA_TRY(cudaMemsetAsync(t->g.mask, 0, (size_t)t->g.mask_stride, t->stream));
    }
    launch_shi_tomasi(e, 1, t->stream, &t->launches);
    GlueParams g1 = t->g;
    g1.n_streams = 1;
    double zero = 0.0;
    VS_CUDA_TRY(cudaMemcpyAsync(t->d_cur_time, &zero, sizeof(double), cudaMemcpyHostToDevice, t->stream));
    g1.cur_time = t->d_cur_time;
    launch_select_finish(g1, t->stream, &t->launches);
    int hdr[kHdrInts];
    VS_CUDA_TRY(cudaMemcpyAsync(hdr, g1.r_hdr, sizeof(hdr), cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
    const int nn = hdr[H_NNEW];
    *n_out = nn;
    if (nn > 0) VS_CUDA_TRY(cudaMemcpyAsync(corners_xy, g1.r_pts + n_kept, sizeof(float) * 2 * nn, cudaMemcpyDeviceToHost, t->stream));
    if (mask_out) VS_CUDA_TRY(cudaMemcpy2DAsync(mask_out, t->W, g1.mask, g1.mask_pitch, t->W, t->H, cudaMemcpyDeviceToHost, t->stream));
    return stage_end(t);
}

int vs_stage_undistort(vs_tracker* t, const float* pts_xy, int32_t n, float* out_xy) {
    int rc = stage_begin(t);
    if (rc != VS_OK) return rc;
    if (n < 0 || (n > 0 && (!pts_xy || !out_xy))) { set_error("bad argument"); return VS_ERR_INVALID; }
    const int64_t room = (int64_t)t->B * t->cap;
    for (int64_t off = 0; off < n; off += room) {
        const int m = (int)((n - off) < room ? (n - off) : room);
        VS_CUDA_TRY(cudaMemcpyAsync(t->g.w_cur, pts_xy + 2 * off, sizeof(float) * 2 * m, cudaMemcpyHostToDevice, t->stream));
        launch_undistort(t->g.cam, t->g.w_cur, m, t->g.w_next, t->stream);
        t->launches += 1;
        VS_CUDA_TRY(cudaMemcpyAsync(out_xy + 2 * off, t->g.w_next, sizeof(float) * 2 * m, cudaMemcpyDeviceToHost, t->stream));
        VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
    }
    return stage_end(t);
}

int vs_stage_set_mask(vs_tracker* t, const float* pts_xy, const int32_t* track_cnt, int32_t n, int32_t* keep_idx,
                      int32_t* n_keep) {
    int rc = stage_begin(t);
    if (rc != VS_OK) return rc;
    if (n < 0 || !n_keep || (n > 0 && (!pts_xy || !track_cnt || !keep_idx))) { set_error("bad argument"); return VS_ERR_INVALID; }
    if (n > t->cap) { set_error("n above capacity"); return VS_ERR_CAPACITY; }
    GlueParams g1 = t->g;
    g1.n_streams = 1;
    std::vector<int> idx(n);
    for (int i = 0; i < n; ++i) idx[i] = i;
    if (n > 0) {
        VS_CUDA_TRY(cudaMemcpyAsync(g1.w_next, pts_xy, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, t->stream));
        VS_CUDA_TRY(cudaMemcpyAsync(g1.w_cnt, track_cnt, sizeof(int) * n, cudaMemcpyHostToDevice, t->stream));
        VS_CUDA_TRY(cudaMemcpyAsync(g1.w_ids, idx.data(), sizeof(int) * n, cudaMemcpyHostToDevice, t->stream));
    }
    VS_CUDA_TRY(cudaMemcpyAsync(g1.wn, &n, sizeof(int), cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(g1.ransac_flag, 0, sizeof(int), t->stream));
    launch_set_mask(g1, t->stream, &t->launches);
    int kn = 0;
    VS_CUDA_TRY(cudaMemcpyAsync(&kn, g1.kn, sizeof(int), cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
    *n_keep = kn;
    if (kn > 0) VS_CUDA_TRY(cudaMemcpyAsync(keep_idx, g1.k_ids, sizeof(int) * kn, cudaMemcpyDeviceToHost, t->stream));
    return stage_end(t);
}

}  // extern "C"
