/*
 * vs_tracker.h — C ABI of the B200-native VINS-Mono visual front-end.
 *
 * Drop-in boundary for the hot path of chansoopark98/Mobile-SLAM:
 *     frontend::FeatureTracker            include/frontend/feature_tracker.h:25-76
 *     FeatureTracker::detectAndTrack      src/frontend/feature_tracker.cpp:97-201
 * The reference has no FFI of its own (its boundary is that C++ class, read through public
 * members by src/utility/measurement_processor.cpp:191-231 and src/vio_engine.cpp:210-239), so
 * the entry points below are what a header-compatible adapter class binds — see
 * include/vs_feature_tracker.hpp (C++ adapter) and INTEGRATION.md (reference-side stub).
 *
 * Plain C: pointers and sizes only, no OpenCV / Eigen / STL / torch types.
 * All functions return 0 on success or a negative vs_status; vs_last_error() gives the text.
 * Calls on one handle must be serialised by the caller (same rule as the reference class).
 * There is NO CPU fallback: every entry point that computes runs CUDA kernels on the
 * handle's device and fails with VS_ERR_CUDA if that is impossible.
 */
#ifndef VS_TRACKER_H_
#define VS_TRACKER_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VS_ABI_VERSION 1

#if defined(__GNUC__)
#define VS_API __attribute__((visibility("default")))
#else
#define VS_API
#endif

struct vs_tracker;              /* opaque handle */
typedef struct vs_tracker vs_tracker;

typedef enum vs_status {
    VS_OK = 0,
    VS_ERR_INVALID = -1,   /* bad argument / config           */
    VS_ERR_CUDA = -2,      /* CUDA runtime failure            */
    VS_ERR_NOMEM = -3,     /* host or device allocation       */
    VS_ERR_CAPACITY = -4,  /* max_cnt above handle capacity   */
    VS_ERR_STATE = -5      /* call not valid in this state    */
} vs_status;

/* common::camera_models::Camera::ModelType, include/common/camera_models/Camera.h:15 */
typedef enum vs_camera_model {
    VS_CAM_KANNALA_BRANDT = 0,
    VS_CAM_MEI = 1,
    VS_CAM_PINHOLE = 2,
    VS_CAM_SCARAMUZZA = 3
} vs_camera_model;

/* Replaces FeatureTracker::readIntrinsicParameter / setIntrinsicParameter
 * (feature_tracker.cpp:318-336) and the m_camera object they build. */
typedef struct vs_camera {
    int32_t model_type;     /* vs_camera_model */
    int32_t width, height;
    double fx, fy, cx, cy;  /* pinhole fx..cy | KB mu,mv,u0,v0 | MEI gamma1,gamma2,u0,v0 | OCAM -, -, center_x, center_y */
    double d[4];            /* pinhole/MEI: k1,k2,p1,p2 | KB: k2,k3,k4,k5 */
    double xi;              /* MEI mirror parameter */
    double poly[5];         /* SCARAMUZZA forward polynomial (SCARAMUZZA_POLY_SIZE = 5) */
    double aff_c, aff_d, aff_e; /* SCARAMUZZA affine C, D, E */
} vs_camera;

/* Replaces the fields of the global utility::g_config the tracker reads on every call
 * (include/utility/config.h:18-56), captured per handle instead of per process. */
typedef struct vs_tracker_config {
    int32_t struct_size;    /* = sizeof(vs_tracker_config) */
    int32_t device;         /* CUDA device ordinal */
    int32_t n_streams;      /* independent camera streams tracked by this handle (>=1) */
    int32_t width, height;  /* g_config.camera.col / row */
    double focal_length;    /* g_config.camera.focal_length (virtual pinhole for F-matrix RANSAC) */
    int32_t max_cnt;        /* feature_tracker.max_cnt  (150) */
    int32_t min_dist;       /* feature_tracker.min_dist (30)  */
    double f_threshold;     /* feature_tracker.f_threshold (1.0) */
    double f_threshold_edge_factor; /* (0 = off) */
    int32_t equalize;       /* CLAHE(3.0, 8x8) on/off (1) */
    int32_t fisheye;        /* circular/auto fisheye mask on/off (0) */
    int32_t lk_window_size;     /* 21 */
    int32_t lk_pyramid_levels;  /* = OpenCV maxLevel: levels 0..N (3) */
    int32_t lk_iterations;      /* 30 */
    double lk_eps;              /* 0.01 */
    int32_t fb_check;       /* 0 = reference semantics (forward LK only); 1 = add backward consistency check */
    double fb_threshold;    /* px, used when fb_check */
    int32_t ransac_mode;    /* vs_ransac_mode */
    int32_t capacity;       /* max features per stream the handle can ever hold (0 = max(max_cnt, 256)) */
    const uint8_t* fisheye_mask_image; /* optional H x W mask (255 = keep), replaces the auto circle; may be NULL */
    int32_t fisheye_mask_stride;
    int32_t first_feature_id;   /* initial value of the per-stream n_id counter (0) */
    int32_t pipeline_lanes;     /* >1: split the handle's streams into that many lanes, each on its own CUDA stream, so
                                   the host->device copy of one lane overlaps the kernels of another (0/1 = one lane) */
    vs_camera camera;
} vs_tracker_config;

typedef enum vs_ransac_mode {
    VS_RANSAC_OFF = 0,      /* skip rejectWithFundamentalMatrix (stage isolation / latency mode) */
    VS_RANSAC_DEVICE = 1,   /* built-in 7-point RANSAC evaluated on the GPU */
    VS_RANSAC_CALLBACK = 2  /* host callback (e.g. cv::findFundamentalMat in the adapter) */
} vs_ransac_mode;

/* Host-side outlier rejection hook (feature_tracker.cpp:203-307).  Points are the float32
 * virtual-pinhole coordinates the reference hands to cv::findFundamentalMat; next_px are the raw
 * pixel positions (for the edge-recovery rule).  Fill status[i] in {0,1}. Return 0 on success. */
typedef int (*vs_reject_fn)(void* user, int32_t stream, int32_t n, const float* un_cur_xy,
                            const float* un_next_xy, const float* next_px_xy, uint8_t* status);

/* Per-stream view of the tracker's public vectors after a frame
 * (cur_pts, cur_undistorted_pts, pts_velocity, ids, track_cnt — feature_tracker.h:53-57).
 * Pointers reference pinned host memory owned by the handle, valid until the next track call. */
typedef struct vs_track_result {
    int32_t n;                /* number of features, all index-aligned below */
    int32_t n_new;            /* the last n_new entries were detected this frame (reference: ids == -1 until updateID) */
    const int32_t* ids;
    const int32_t* track_cnt;
    const float* cur_pts;     /* n x 2, distorted pixel coordinates */
    const float* undist;      /* n x 2, liftProjective x/z, y/z */
    const float* velocity;    /* n x 2 */
    /* per-frame counters */
    int32_t n_lk_in, n_lk_ok, n_ransac_in, n_ransac_out, n_mask_kept, n_candidates;
    int32_t ransac_ran;
    float rms_disp;
} vs_track_result;

VS_API const char* vs_last_error(void);
VS_API int vs_abi_version(void);

/* FeatureTracker() + readIntrinsicParameter/setIntrinsicParameter */
VS_API int vs_tracker_create(const vs_tracker_config* cfg, vs_tracker_config* effective_or_null, struct vs_tracker** out);
VS_API void vs_tracker_destroy(struct vs_tracker* t);
VS_API void vs_tracker_default_config(vs_tracker_config* cfg);

/* detectAndTrack(img, t) + the caller's updateID loop, one stream (n_streams must be 1). */
VS_API int vs_tracker_track(struct vs_tracker* t, const uint8_t* gray, int32_t stride, double time_s,
                     vs_track_result* out);
/* One frame for each of the handle's streams; frames[s] is an H x stride host image. */
VS_API int vs_tracker_track_batch(struct vs_tracker* t, const uint8_t* const* frames, int32_t stride,
                           const double* time_s, vs_track_result* out /* [n_streams] */);
/* Asynchronous form of vs_tracker_track_batch: submit enqueues upload + kernels + result download and returns at once;
 * wait blocks until the OLDEST submitted batch is done and describes it in out[n_streams].  Up to two batches may be
 * in flight, so the host->device copy of batch k+1 overlaps the kernels of batch k.  The frame buffers of a batch must
 * stay valid until its wait returns; its results stay valid until the second wait after it. */
VS_API int vs_tracker_submit_batch(struct vs_tracker* t, const uint8_t* const* frames, int32_t stride,
                                   const double* time_s);
VS_API int vs_tracker_wait(struct vs_tracker* t, vs_track_result* out /* [n_streams] */);
/* Same, frames already resident in device memory: n_streams images, each H rows of `stride` bytes,
 * `image_stride_bytes` apart.  results_to_host = 0 leaves results on the device (fetch later). */
VS_API int vs_tracker_track_batch_device(struct vs_tracker* t, const uint8_t* d_frames, int32_t stride,
                                  int64_t image_stride_bytes, const double* time_s,
                                  int32_t results_to_host, vs_track_result* out_or_null);
VS_API int vs_tracker_fetch_results(struct vs_tracker* t, vs_track_result* out /* [n_streams] */);

/* Runtime mutators, VIOEngine::setMobileParams/setFThreshold/setTrackingParams (vio_engine.cpp:354-389).
 * Negative / zero values leave a field unchanged, exactly like the reference's guards. */
VS_API int vs_tracker_set_params(struct vs_tracker* t, int32_t max_cnt, int32_t min_dist, double f_threshold,
                          int32_t lk_window, int32_t lk_pyramid, int32_t lk_iterations, double lk_eps,
                          double f_edge_factor);
/* The remaining g_config fields the reference re-reads on every call (feature_tracker.cpp:39-46,52-55,102,214-222):
 * equalize (CLAHE on/off), fisheye (mask on/off; the mask = fisheye_mask_image if given, else the auto circle) and
 * camera.focal_length.  Negative equalize / fisheye, focal_length <= 0 and a NULL image leave that field unchanged.
 * Tracking state is kept.  Not valid while a submitted batch is in flight. */
VS_API int vs_tracker_set_frontend_flags(struct vs_tracker* t, int32_t equalize, int32_t fisheye, double focal_length,
                                         const uint8_t* fisheye_mask_image, int32_t fisheye_mask_stride);
VS_API int vs_tracker_set_reject_callback(struct vs_tracker* t, vs_reject_fn fn, void* user);
VS_API int vs_tracker_reset(struct vs_tracker* t, int32_t stream /* -1 = all */);
/* number of CUDA kernels this handle has launched so far (bench.py "gpu_launches") */
/* Debug aid: the newest frame's pyramid block of one stream exactly as the LK kernel sees it (padded levels).
 * layout[4 l .. 4 l + 3] = byte offset of pixel (0, 0), pitch, width, height of level l (l < 6); layout[24] = block size
 * in bytes, layout[25] / layout[26] = border columns / rows stored around every level.  out may be NULL (layout only). */
VS_API int vs_tracker_debug_read_pyramid(struct vs_tracker* t, int32_t stream, uint8_t* out, int64_t cap_bytes,
                                         int64_t* layout /* [27] */, int32_t* n_levels);
VS_API int64_t vs_tracker_launch_count(const struct vs_tracker* t);
/* Debug aid: a handle created with VS_DEBUG_GUARD=1 in the environment brackets every device buffer with 4 KiB guard
 * bands; this synchronises the device and returns how many guard bytes have been overwritten since creation
 * (0 = clean, -1 = the handle was created without guards, -2 = CUDA error). */
VS_API int64_t vs_tracker_debug_guard_check(struct vs_tracker* t);
/* CUDA-event time (ms) of the device work of the last track call */
VS_API float vs_tracker_last_device_ms(const struct vs_tracker* t);

/* ---- stage entry points (same kernels as the frame path, for teacher-forced parity tests and
 *      per-kernel roofline timing).  All buffers are host memory, images are tightly packed. ---- */
VS_API int vs_stage_clahe(struct vs_tracker* t, const uint8_t* img, uint8_t* out);
/* levels written back to back, tightly packed; sizes[2*l], sizes[2*l+1] = width, height */
VS_API int vs_stage_pyramid(struct vs_tracker* t, const uint8_t* level0, uint8_t* out_levels, int32_t* sizes,
                     int32_t* n_levels);
VS_API int vs_stage_lk(struct vs_tracker* t, const uint8_t* prev_level0, const uint8_t* next_level0,
                const float* pts_xy, int32_t n, float* next_xy, uint8_t* status /* raw LK status */);
VS_API int vs_stage_min_eig(struct vs_tracker* t, const uint8_t* img, float* eig);
/* mask = base (fisheye/all) minus min_dist discs around kept_xy; returns corners in acceptance order */
VS_API int vs_stage_good_features(struct vs_tracker* t, const uint8_t* img, const float* kept_xy, int32_t n_kept,
                           int32_t max_corners, float* corners_xy, int32_t* n_out, uint8_t* mask_out_or_null);
VS_API int vs_stage_undistort(struct vs_tracker* t, const float* pts_xy, int32_t n, float* out_xy);
/* setMask order + greedy keep (feature_tracker.cpp:62-86): keep_idx = indices into the input, in kept order */
VS_API int vs_stage_set_mask(struct vs_tracker* t, const float* pts_xy, const int32_t* track_cnt, int32_t n,
                      int32_t* keep_idx, int32_t* n_keep);
/* F-matrix RANSAC on virtual-pinhole points (device implementation) */
VS_API int vs_stage_ransac(struct vs_tracker* t, const float* a_xy, const float* b_xy, int32_t n, uint8_t* status);
/* Flags for a GIVEN fundamental matrix f9 (row-major, x2^T F x1 = 0): base_status == NULL -> inlier test (both epipolar
 * distances <= f_threshold) + the reference's edge recovery (feature_tracker.cpp:236-285, edge_factor > 0); base_status
 * != NULL -> those flags (e.g. cv::findFundamentalMat's own mask) + only the edge-recovery rule.  Teacher-forced parity. */
VS_API int vs_stage_ransac_mark(struct vs_tracker* t, const double* f9, const float* a_xy, const float* b_xy,
                                const float* next_px_xy, int32_t n, double edge_factor, const uint8_t* base_status,
                                uint8_t* status);
/* ---- allocation-free entry point ----
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * CLAHE(clipLimit 3.0, 8 x 8 tiles) of n_images 8-bit DEVICE images, the same two kernels as vs_stage_clahe and the frame
 * path (bit-identical output, == cv::CLAHE::apply).  d_src / d_dst: device memory of the CURRENT device, image i at
 * d_* + i * *_image_stride, rows *_pitch bytes apart; pointers, pitches and image strides multiples of 16, pitches >=
 * width, 32 <= width, height <= 16384; d_src and d_dst must not overlap.  cuda_stream: a cudaStream_t (NULL = the
 * default stream).  The call enqueues two kernels on that stream and returns without synchronising; it needs no
 * tracker handle, calls no cudaMalloc / cudaFree / allocating library routine and creates no stream or event.
 * Limit: n_images <= VS_CLAHE_NO_ALLOC_MAX_IMAGES, else VS_ERR_CAPACITY (split the batch into calls on ONE stream).
 * Its scratch — the 64 tile look-up tables of each image, 16 KiB per image — is a static device array inside the
 * library: every GPU that loads the library holds VS_CLAHE_NO_ALLOC_MAX_IMAGES x 16 KiB = 1 MiB until the process
 * exits; ALL calls on a device share it, so calls that overlap on the GPU (different streams, concurrent launches of one
 * captured graph) RACE on it — calls on one stream are ordered and safe; it keeps the previous call's contents, which is
 * harmless: every table entry a call reads is written by that call first. */
#define VS_CLAHE_NO_ALLOC_MAX_IMAGES 64
VS_API int vs_clahe_no_alloc(const uint8_t* d_src, int32_t src_pitch, int64_t src_image_stride, uint8_t* d_dst,
                             int32_t dst_pitch, int64_t dst_image_stride, int32_t width, int32_t height,
                             int32_t n_images, void* cuda_stream);
/* ---- measurement hooks (bench.py) ---- */
/* The handle's CUDA stream (a cudaStream_t): lets a caller record its own events on the stream the
 * kernels are launched on, or order its own work after a track call. */
VS_API void* vs_tracker_cuda_stream(const struct vs_tracker* t);
/* Per-kernel CUDA-event timing of the frame path.  on != 0 brackets every kernel of each following
 * track call with events on the handle's stream and accumulates the elapsed times (it adds a
 * stream synchronise per call, so it is for profiling passes, not the timed throughput region);
 * switching it on also clears the accumulators. */
VS_API int vs_tracker_set_profiling(struct vs_tracker* t, int32_t on);
/* The records the tracker's two callers build from its public vectors, for one stream's result:
 *   MeasurementProcessor::extractImageFeatures (src/utility/measurement_processor.cpp:210-231): ray (x, y, 1) + channels
 *   id, u, v, vx, vy;  VIOEngine::processFrame (src/vio_engine.cpp:219-239): common::ImageData id -> [x, y, 1, u, v, vx, vy].
 * Only features with track_cnt > 1 are published, in array order.  ids_out[max_n], rec_out[max_n][7] doubles
 * (x = undist.x, y = undist.y, 1, u = cur_pts.x, v = cur_pts.y, vx, vy).  Host-side, needs no GPU.  *n_out = records
 * written; VS_ERR_CAPACITY if max_n is too small (nothing beyond max_n is written). */
VS_API int vs_pack_image_features(const vs_track_result* r, int32_t max_n, int32_t* ids_out, double* rec_out, int32_t* n_out);

/* A copy of the handle's current configuration (after vs_tracker_set_params / vs_tracker_set_frontend_flags);
 * fisheye_mask_image is always NULL in the copy. */
VS_API int vs_tracker_get_config(const struct vs_tracker* t, vs_tracker_config* out);

/* ---- frame ingest: the step before the path -------------------------------------------------------------------
 * MeasurementProcessor::extractImageFeatures (src/utility/measurement_processor.cpp:178-231) reads every frame with
 * cv::imread(path, IMREAD_GRAYSCALE), tracks rowRange(0, g_config.camera.row) of it, runs the updateID loop and publishes
 * the track_cnt > 1 features; VIOEngine::processFrame (src/vio_engine.cpp:180) wraps a raw grey pointer that is only
 * valid during the call.  Host-side (no GPU needed for the two decode functions). */
/* PNG (all colour types / bit depths, non-interlaced) or binary PGM -> 8-bit grey exactly as cv::imread(IMREAD_GRAYSCALE)
 * returns it.  Writes min(image height, max_h) rows of out_stride bytes; *w, *h = the file's own size.  VS_ERR_INVALID
 * ("Cannot load image: ...") for unreadable / unsupported / wider-than-max_w files. */
VS_API int vs_image_decode_gray(const uint8_t* file_bytes, int64_t n_bytes, uint8_t* out, int32_t out_stride,
                                int32_t max_w, int32_t max_h, int32_t* w, int32_t* h);
VS_API int vs_image_read_gray(const char* path, uint8_t* out, int32_t out_stride, int32_t max_w, int32_t max_h,
                              int32_t* w, int32_t* h);
/* Pipeline around one handle: decode threads fill a ring of three pinned host slots, one coordinator thread drives
 * vs_tracker_submit_batch / vs_tracker_wait, so batch k+2 is decoded while batch k+1 uploads and batch k is in the
 * kernels.  submit_* return at once (VS_ERR_STATE when three batches are outstanding); vs_ingest_wait blocks for the
 * OLDEST batch and describes it in out[n_streams] (arrays owned by the pipeline, valid until the next vs_ingest_wait).
 * A file that cannot be decoded fails its batch's wait with VS_ERR_INVALID and the file name in vs_last_error().
 * While a pipeline exists the handle must not be used directly.  Not available with VS_RANSAC_CALLBACK. */
struct vs_ingest;
typedef struct vs_ingest vs_ingest;
VS_API int vs_ingest_create(struct vs_tracker* t, int32_t decode_threads, struct vs_ingest** out);
VS_API int vs_ingest_submit_files(struct vs_ingest* g, const char* const* paths /* [n_streams] */, const double* time_s);
VS_API int vs_ingest_submit_raw(struct vs_ingest* g, const uint8_t* const* frames, int32_t stride, const double* time_s);
VS_API int vs_ingest_wait(struct vs_ingest* g, vs_track_result* out /* [n_streams] */);
VS_API void vs_ingest_destroy(struct vs_ingest* g);

/* ---- device feature bank: frontend::FeatureManager for n_streams independent streams -----------------
 * Replaces, per stream, the class of /root/reference/include/frontend/feature_manager.h:70-92 whose methods live in
 * /root/reference/src/frontend/feature_manager.cpp (cited per entry point).  Feature records are rows of 7 doubles
 * [x, y, z, u, v, vx, vy] (common/image_frame.h:14, the layout vs_pack_image_features writes).  WINDOW_SIZE is 10
 * (utility/config.h:11): a feature holds at most VS_BANK_MAX_OBS observations.  Decisions, parallax sums and shifted
 * depths are bit-equal to the reference; triangulated depths agree to rounding (the SVD is a Jacobi iteration here,
 * Eigen::JacobiSVD there).  All calls synchronise the bank's own stream before returning; calls on one handle are
 * serialised by the caller (like the reference's manager, which lives on the estimator's thread).  A stream whose bank
 * overflowed (VS_ERR_CAPACITY from an add) holds an incomplete frame: every later add on the handle reports it again
 * until vs_bank_clear resets that stream. */
#define VS_BANK_WINDOW_SIZE 10
#define VS_BANK_MAX_OBS (VS_BANK_WINDOW_SIZE + 1)
struct vs_bank;
VS_API int vs_bank_create(int32_t device, int32_t n_streams, int32_t capacity /* features per stream */, struct vs_bank** out);
VS_API void vs_bank_destroy(struct vs_bank* b);
/* clearState, feature_manager.cpp:11-13; stream -1 = all */
VS_API int vs_bank_clear(struct vs_bank* b, int32_t stream);
/* addFeatureAndCheckParallax, feature_manager.cpp:27-73 (+ compensatedParallax2 :264-294), every stream in one launch.
 * frame_count[s], n[s] records ids[s][max_n] / rec[s][max_n][7] (unique ids, any order: the std::map order of
 * common::ImageData is applied).  is_key[s] = the method's return value; optional outputs: last_track_num_, and the
 * parallax sum / count the decision was taken on.  VS_ERR_CAPACITY when a stream's bank is full. */
VS_API int vs_bank_add_frame(struct vs_bank* b, const int32_t* frame_count, const int32_t* n, const int32_t* ids,
                             const double* rec, int32_t max_n, double min_parallax, double focal_length, int32_t* is_key,
                             int32_t* last_track_num, double* parallax_sum, int32_t* parallax_num);
/* The same, fed from the DEVICE-resident results of the tracker handle's last frame: the publish loop (track_cnt > 1,
 * measurement_processor.cpp:210-231, vio_engine.cpp:219-239) runs on the device and only the decisions cross PCIe.
 * The bank must have the handle's stream count and device. */
VS_API int vs_bank_add_frame_from_tracker(struct vs_bank* b, struct vs_tracker* t, const int32_t* frame_count,
                                          double min_parallax, double focal_length, int32_t* is_key, int32_t* last_track_num);
/* Window maintenance, op[s]: 0 nothing, 1 removeBack (feature_manager.cpp:237-249), 2 removeFront(frame_count[s])
 * (:251-266), 3 removeBackShiftDepth (:208-235; rp[s] = marg_R (9, row-major) marg_P (3) new_R (9) new_P (3)),
 * 4 removeFailures (:112-118).  init_depth = g_config.estimator.init_depth. */
VS_API int vs_bank_slide(struct vs_bank* b, const int32_t* op, const int32_t* frame_count_or_null,
                         const double* rp_or_null /* [n_streams][24] */, double init_depth);
/* triangulateAcrossAllViews, feature_manager.cpp:142-194: poses[s][VS_BANK_MAX_OBS][12] = sliding_window[i].R (row-major) | .P;
 * active_or_null[s] == 0 leaves stream s alone (its estimator has no full window yet) */
VS_API int vs_bank_triangulate(struct vs_bank* b, const double* poses, const double* t_ic /* [3] */,
                               const double* r_ic /* [9] row-major */, double init_depth, const int32_t* active_or_null);
/* getDepthVector (:130-140) of one stream, *n = getFeatureCount (:15-25); out may be null to ask for the count */
VS_API int vs_bank_get_depth_vector(struct vs_bank* b, int32_t stream, int32_t max_n, double* out, int32_t* n);
/* setDepth (:94-110, clear = 0) / clearDepth (:120-128, clear = 1) */
VS_API int vs_bank_set_depth(struct vs_bank* b, int32_t stream, const double* x, int32_t n, int32_t clear);
/* getCorresponding (:75-92): pairs[k] = ray in frame_l (3) | ray in frame_r (3), list order; pairs may be null */
VS_API int vs_bank_get_corresponding(struct vs_bank* b, int32_t stream, int32_t frame_l, int32_t frame_r, int32_t max_n,
                                     double* pairs, int32_t* n);
/* Debug aid like vs_tracker_debug_guard_check: with VS_DEBUG_GUARD=1 in the environment at vs_bank_create every device
 * block of the bank lies between two guard bands; returns the number of guard bytes that changed (0 = no kernel stored
 * outside its buffers), -1 when the bank was created without guards. */
VS_API int64_t vs_bank_debug_guard_check(struct vs_bank* b);
/* feature_bank_ of one stream in list order (any output may be null); obs[max_n][VS_BANK_MAX_OBS][7] */
VS_API int vs_bank_export(struct vs_bank* b, int32_t stream, int32_t max_n, int32_t* ids, int32_t* start_frame,
                          int32_t* n_obs, double* depth, int32_t* solve_flag, double* obs, int32_t* n);

/* total_ms[i] / launches[i] for kernel slot i < *n_out <= max_n; names from vs_profile_kernel_name */
VS_API int vs_tracker_get_profile(const struct vs_tracker* t, int32_t max_n, float* total_ms, int32_t* launches,
                                  int32_t* n_out);
VS_API const char* vs_profile_kernel_name(int32_t slot);

#ifdef __cplusplus
}
#endif
#endif /* VS_TRACKER_H_ */
