// vs_feature_tracker.hpp — header-only C++ host layer over the C ABI of include/vs_tracker.h.
//
// `frontend::FeatureTracker` below is member-for-member compatible with the class the reference's callers use
// (/root/reference/include/frontend/feature_tracker.h:25-76): same method names
// (detectAndTrack / readImage, updateID, setIntrinsicParameter, readIntrinsicParameter) and the same public vectors
// (cur_pts, cur_undistorted_pts, pts_velocity, ids, track_cnt), which src/utility/measurement_processor.cpp:191-231
// and src/vio_engine.cpp:210-239 read directly.  All computation happens in libvs_tracker.so (CUDA, sm_100a);
// this file only marshals buffers.  There is no CPU fallback: construction of the handle throws if CUDA is missing.
//
// Build modes
//   -DVS_WITH_OPENCV          point type = cv::Point2f, image type = cv::Mat, and the F-matrix outlier rejection of
//                             feature_tracker.cpp:203-307 runs through cv::findFundamentalMat (VS_RANSAC_CALLBACK),
//                             i.e. bit-for-bit the reference's own call.
//   (default, no OpenCV)      vs::Point2f / vs::ImageView stand-ins with the same field names (x, y / data, cols, rows,
//                             step); rejection = the built-in device RANSAC (VS_RANSAC_DEVICE, default) or VS_RANSAC_OFF as configured.
//   -DVS_WITH_REFERENCE_CONFIG  read utility::g_config (include/utility/config.h:18-56) on every call like the
//                             reference does; otherwise use the `config` member.
#ifndef VS_FEATURE_TRACKER_HPP_
#define VS_FEATURE_TRACKER_HPP_

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "vs_tracker.h"

#ifdef VS_WITH_OPENCV
#include <opencv2/opencv.hpp>
#endif
#ifdef VS_WITH_REFERENCE_CONFIG
#include "utility/config.h"
#endif

namespace vs {

#ifdef VS_WITH_OPENCV
using Point2f = cv::Point2f;
#else
struct Point2f {
    float x = 0.f, y = 0.f;
    Point2f() = default;
    Point2f(float x_, float y_) : x(x_), y(y_) {}
};
// the fields of cv::Mat the tracker touches (CV_8UC1)
struct ImageView {
    const uint8_t* data = nullptr;
    int cols = 0, rows = 0;
    size_t step = 0;     // bytes per row
    ImageView() = default;
    ImageView(const uint8_t* d, int w, int h, size_t s = 0) : data(d), cols(w), rows(h), step(s ? s : (size_t)w) {}
};
#endif

// the fields of utility::g_config the tracker reads (include/utility/config.h:18-56), per instance
struct FrontendConfig {
    int row = 480, col = 752;
    double focal_length = 460.0;
    int max_cnt = 150, min_dist = 30;
    double f_threshold = 1.0, f_threshold_edge_factor = 0.0;
    int equalize = 1, fisheye = 0;
    int lk_window_size = 21, lk_pyramid_levels = 3, lk_iterations = 30;
    double lk_eps = 0.01;
    int device = 0;
#ifdef VS_WITH_OPENCV
    int ransac_mode = VS_RANSAC_CALLBACK;
#else
    int ransac_mode = VS_RANSAC_DEVICE;   // the reference always rejects outliers; without OpenCV the built-in device RANSAC does
#endif
};

class Error : public std::runtime_error {
public:
    Error(int code, const char* what) : std::runtime_error(std::string("vs_tracker error ") + std::to_string(code) + ": " + what) {}
};
inline void check(int rc) { if (rc != VS_OK) throw Error(rc, vs_last_error()); }

// CameraFactory::generateCameraFromYamlFile (src/common/camera_models/CameraFactory.cc:87-150) for the OpenCV
// FileStorage YAML the reference ships (config/*.yaml): flat `key: value` pairs under one level of sections.
inline bool camera_from_yaml(const std::string& path, vs_camera* cam) {
    std::ifstream f(path);
    if (!f) return false;
    std::map<std::string, std::string> kv;
    std::string line, section;
    while (std::getline(f, line)) {
        const size_t hash = line.find('#');
        if (hash != std::string::npos) line.erase(hash);
        if (line.empty() || line[0] == '%' || line.find(':') == std::string::npos) continue;
        const size_t indent = line.find_first_not_of(" \t");
        if (indent == std::string::npos) continue;
        const size_t colon = line.find(':');
        auto trim = [](std::string s) {
            const size_t a = s.find_first_not_of(" \t\"\r"), b = s.find_last_not_of(" \t\"\r");
            return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
        };
        const std::string key = trim(line.substr(0, colon)), val = trim(line.substr(colon + 1));
        if (indent == 0) {
            if (val.empty() || val.rfind("!!", 0) == 0) section = key; else { section.clear(); kv[key] = val; }
        } else if (!section.empty()) {
            kv[section + "." + key] = val;
        }
    }
    auto num = [&](const std::string& k) { auto it = kv.find(k); return it == kv.end() ? 0.0 : std::atof(it->second.c_str()); };
    std::memset(cam, 0, sizeof(*cam));
    std::string model = kv.count("model_type") ? kv["model_type"] : "MEI";     // factory default (CameraFactory.cc:94)
    for (auto& c : model) c = (char)std::toupper((unsigned char)c);
    cam->width = (int)num("image_width");
    cam->height = (int)num("image_height");
    cam->aff_c = 1.0;
    if (model == "PINHOLE") {
        cam->model_type = VS_CAM_PINHOLE;
        cam->fx = num("projection_parameters.fx"); cam->fy = num("projection_parameters.fy");
        cam->cx = num("projection_parameters.cx"); cam->cy = num("projection_parameters.cy");
        cam->d[0] = num("distortion_parameters.k1"); cam->d[1] = num("distortion_parameters.k2");
        cam->d[2] = num("distortion_parameters.p1"); cam->d[3] = num("distortion_parameters.p2");
    } else if (model == "KANNALA_BRANDT") {
        cam->model_type = VS_CAM_KANNALA_BRANDT;
        cam->fx = num("projection_parameters.mu"); cam->fy = num("projection_parameters.mv");
        cam->cx = num("projection_parameters.u0"); cam->cy = num("projection_parameters.v0");
        cam->d[0] = num("projection_parameters.k2"); cam->d[1] = num("projection_parameters.k3");
        cam->d[2] = num("projection_parameters.k4"); cam->d[3] = num("projection_parameters.k5");
    } else if (model == "MEI") {
        cam->model_type = VS_CAM_MEI;
        cam->xi = num("mirror_parameters.xi");
        cam->fx = num("projection_parameters.gamma1"); cam->fy = num("projection_parameters.gamma2");
        cam->cx = num("projection_parameters.u0"); cam->cy = num("projection_parameters.v0");
        cam->d[0] = num("distortion_parameters.k1"); cam->d[1] = num("distortion_parameters.k2");
        cam->d[2] = num("distortion_parameters.p1"); cam->d[3] = num("distortion_parameters.p2");
    } else if (model == "SCARAMUZZA") {
        cam->model_type = VS_CAM_SCARAMUZZA;
        cam->fx = cam->fy = 1.0;
        cam->cx = num("affine_parameters.cx"); cam->cy = num("affine_parameters.cy");
        cam->aff_c = kv.count("affine_parameters.ac") ? num("affine_parameters.ac") : 1.0;
        cam->aff_d = num("affine_parameters.ad"); cam->aff_e = num("affine_parameters.ae");
        for (int i = 0; i < 5; ++i) cam->poly[i] = num("poly_parameters.p" + std::to_string(i));
    } else {
        return false;        // CameraFactory returns CameraPtr() for an unknown model
    }
    return true;
}

}  // namespace vs

namespace frontend {

class FeatureTracker {
public:
#ifdef VS_WITH_OPENCV
    using Image = cv::Mat;
#else
    using Image = vs::ImageView;
#endif
    using Point2f = vs::Point2f;

    FeatureTracker() = default;
    ~FeatureTracker() { if (handle_) vs_tracker_destroy(handle_); }
    FeatureTracker(const FeatureTracker&) = delete;
    FeatureTracker& operator=(const FeatureTracker&) = delete;

    // feature_tracker.cpp:97-201 (one call = CLAHE -> pyramid -> LK -> filter -> RANSAC -> setMask -> Shi-Tomasi ->
    // addPoints -> undistortedPoints, all inside vs_tracker_track)
    void detectAndTrack(const Image& img, double _cur_time) {
        ensure_handle(img.cols, img.rows);
        apply_runtime_params();
        vs_track_result r;
        vs::check(vs_tracker_track(handle_, img.data, (int32_t)row_bytes(img), _cur_time, &r));
        prev_time = cur_time;
        cur_time = _cur_time;
        const size_t n = (size_t)r.n;
        cur_pts.resize(n); cur_undistorted_pts.resize(n); pts_velocity.resize(n);
        ids.resize(n); track_cnt.resize(n); real_ids_.assign(r.ids, r.ids + n);
        for (size_t i = 0; i < n; ++i) {
            cur_pts[i] = Point2f(r.cur_pts[2 * i], r.cur_pts[2 * i + 1]);
            cur_undistorted_pts[i] = Point2f(r.undist[2 * i], r.undist[2 * i + 1]);
            pts_velocity[i] = Point2f(r.velocity[2 * i], r.velocity[2 * i + 1]);
            track_cnt[i] = r.track_cnt[i];
            // the reference leaves freshly detected points at id -1 until the caller's updateID loop (ft:89-95)
            ids[i] = i + (size_t)r.n_new >= n ? -1 : r.ids[i];
        }
        last = r;
    }
    void readImage(const Image& img, double t) { detectAndTrack(img, t); }     // upstream VINS-Mono's name

    // feature_tracker.cpp:309-316
    bool updateID(unsigned int i) {
        if (i < ids.size()) {
            if (ids[i] == -1) { ids[i] = real_ids_[i]; if (real_ids_[i] + 1 > n_id) n_id = real_ids_[i] + 1; }
            return true;
        }
        return false;
    }

    // feature_tracker.cpp:318-320
    void readIntrinsicParameter(const std::string& calib_file) {
        have_camera_ = vs::camera_from_yaml(calib_file, &camera_);
        if (!have_camera_) std::fprintf(stderr, "[vs] cannot build a camera from %s\n", calib_file.c_str());
        reset_handle();
    }
    // feature_tracker.cpp:322-336: anything but PINHOLE becomes the equidistant model; k2..k5 map to k1,k2,p1,p2 for pinhole
    void setIntrinsicParameter(int model_type, int width, int height, double fx, double fy, double cx, double cy,
                               double k2, double k3, double k4, double k5) {
        std::memset(&camera_, 0, sizeof(camera_));
        camera_.model_type = model_type == VS_CAM_PINHOLE ? VS_CAM_PINHOLE : VS_CAM_KANNALA_BRANDT;
        camera_.width = width; camera_.height = height;
        camera_.fx = fx; camera_.fy = fy; camera_.cx = cx; camera_.cy = cy;
        camera_.d[0] = k2; camera_.d[1] = k3; camera_.d[2] = k4; camera_.d[3] = k5;
        camera_.aff_c = 1.0;
        have_camera_ = true;
        reset_handle();
    }

    // common::ImageData the callers assemble (measurement_processor.cpp:212-223, vio_engine.cpp:219-239):
    // id -> [x, y, 1, u, v, vx, vy] for every feature with track_cnt > 1
    std::map<int, std::vector<double>> imageData() const {
        std::map<int, std::vector<double>> out;
        for (size_t j = 0; j < ids.size(); ++j)
            if (track_cnt[j] > 1)
                out[ids[j]] = {cur_undistorted_pts[j].x, cur_undistorted_pts[j].y, 1.0, cur_pts[j].x, cur_pts[j].y,
                               pts_velocity[j].x, pts_velocity[j].y};
        return out;
    }

    // ---- the public data members the reference's callers read ----
    std::vector<Point2f> cur_pts;
    std::vector<Point2f> cur_undistorted_pts;
    std::vector<Point2f> pts_velocity;
    std::vector<int> ids;
    std::vector<int> track_cnt;
    double cur_time = 0.0;
    double prev_time = 0.0;
    static inline int n_id = 0;          // static int FeatureTracker::n_id (ft:10); ids themselves are per stream
    vs::FrontendConfig config;           // stand-in for utility::g_config when VS_WITH_REFERENCE_CONFIG is off
    vs_track_result last{};              // per-frame counters of the last call
    vs_tracker* handle() const { return handle_; }

private:
    static size_t row_bytes(const Image& img) {
#ifdef VS_WITH_OPENCV
        return img.step[0];
#else
        return img.step;
#endif
    }
    vs::FrontendConfig current() const {
        vs::FrontendConfig c = config;
#ifdef VS_WITH_REFERENCE_CONFIG
        const auto& g = utility::g_config;
        c.row = (int)g.camera.row; c.col = (int)g.camera.col; c.focal_length = g.camera.focal_length;
        c.max_cnt = g.feature_tracker.max_cnt; c.min_dist = g.feature_tracker.min_dist;
        c.f_threshold = g.feature_tracker.f_threshold; c.f_threshold_edge_factor = g.feature_tracker.f_threshold_edge_factor;
        c.equalize = g.feature_tracker.equalize; c.fisheye = g.feature_tracker.fisheye;
        c.lk_window_size = g.feature_tracker.lk_window_size; c.lk_pyramid_levels = g.feature_tracker.lk_pyramid_levels;
        c.lk_iterations = g.feature_tracker.lk_iterations; c.lk_eps = g.feature_tracker.lk_eps;
#endif
        return c;
    }
    void reset_handle() { if (handle_) { vs_tracker_destroy(handle_); handle_ = nullptr; } }
    void ensure_handle(int cols, int rows) {
        if (handle_) return;
        if (!have_camera_) throw std::runtime_error("FeatureTracker: camera intrinsics not set");
        const vs::FrontendConfig c = current();
        vs_tracker_config tc;
        vs_tracker_default_config(&tc);
        tc.device = c.device;
        tc.n_streams = 1;
        tc.width = cols; tc.height = rows;       // the reference hands rowRange(0, g_config.camera.row) (mp:191)
        tc.focal_length = c.focal_length;
        tc.max_cnt = c.max_cnt; tc.min_dist = c.min_dist;
        tc.f_threshold = c.f_threshold; tc.f_threshold_edge_factor = c.f_threshold_edge_factor;
        tc.equalize = c.equalize; tc.fisheye = c.fisheye;
        tc.lk_window_size = c.lk_window_size; tc.lk_pyramid_levels = c.lk_pyramid_levels;
        tc.lk_iterations = c.lk_iterations; tc.lk_eps = c.lk_eps;
        tc.ransac_mode = c.ransac_mode;
        tc.capacity = c.max_cnt > 1024 ? c.max_cnt : 1024;    // room for setMobileParams to raise max_cnt later
        tc.camera = camera_;
        vs::check(vs_tracker_create(&tc, nullptr, &handle_));
        applied_ = c;
#ifdef VS_WITH_OPENCV
        if (c.ransac_mode == VS_RANSAC_CALLBACK) vs::check(vs_tracker_set_reject_callback(handle_, &FeatureTracker::reject_cv, this));
#endif
    }
    // VIOEngine::setMobileParams / setFThreshold / setTrackingParams mutate g_config between frames (vio_engine.cpp:354-389)
    void apply_runtime_params() {
        const vs::FrontendConfig c = current();
        if (c.max_cnt != applied_.max_cnt || c.min_dist != applied_.min_dist || c.f_threshold != applied_.f_threshold ||
            c.lk_window_size != applied_.lk_window_size || c.lk_pyramid_levels != applied_.lk_pyramid_levels ||
            c.lk_iterations != applied_.lk_iterations || c.lk_eps != applied_.lk_eps ||
            c.f_threshold_edge_factor != applied_.f_threshold_edge_factor) {
            vs::check(vs_tracker_set_params(handle_, c.max_cnt, c.min_dist, c.f_threshold, c.lk_window_size, c.lk_pyramid_levels,
                                            c.lk_iterations, c.lk_eps, c.f_threshold_edge_factor));
            applied_ = c;
        }
    }
#ifdef VS_WITH_OPENCV
    // rejectWithFundamentalMatrix, feature_tracker.cpp:226-299: the lift to the virtual pinhole was done on the GPU,
    // the RANSAC itself is the reference's own OpenCV call, edge recovery as in :236-285
    static int reject_cv(void* user, int32_t, int32_t n, const float* a, const float* b, const float* px, uint8_t* status) {
        auto* self = static_cast<FeatureTracker*>(user);
        const vs::FrontendConfig c = self->current();
        std::vector<cv::Point2f> A(n), B(n);
        for (int i = 0; i < n; ++i) { A[i] = {a[2 * i], a[2 * i + 1]}; B[i] = {b[2 * i], b[2 * i + 1]}; }
        std::vector<uchar> st;
        cv::Mat F = cv::findFundamentalMat(A, B, cv::FM_RANSAC, c.f_threshold, 0.99, st);
        for (int i = 0; i < n; ++i) status[i] = i < (int)st.size() ? st[i] : 0;
        if (c.f_threshold_edge_factor > 0.0 && !F.empty() && F.rows == 3) {
            const double cx = c.col / 2.0, cy = c.row / 2.0, rmax = std::sqrt(cx * cx + cy * cy);
            cv::Mat Fd; F.convertTo(Fd, CV_64F);
            for (int i = 0; i < n; ++i) {
                if (status[i]) continue;
                const double dx = px[2 * i] - cx, dy = px[2 * i + 1] - cy, rr = std::sqrt(dx * dx + dy * dy) / rmax;
                if (rr < 0.3) continue;
                const double thr = c.f_threshold * (1.0 + c.f_threshold_edge_factor * rr * rr);
                const cv::Mat x1 = (cv::Mat_<double>(3, 1) << A[i].x, A[i].y, 1.0), x2 = (cv::Mat_<double>(3, 1) << B[i].x, B[i].y, 1.0);
                const cv::Mat l = Fd * x1;
                const double num = std::abs(x2.dot(l)), den = std::sqrt(l.at<double>(0) * l.at<double>(0) + l.at<double>(1) * l.at<double>(1));
                if (den > 1e-10 && num / den < thr) status[i] = 1;
            }
        }
        return 0;
    }
#endif
    vs_tracker* handle_ = nullptr;
    vs_camera camera_{};
    bool have_camera_ = false;
    vs::FrontendConfig applied_{};
    std::vector<int> real_ids_;
};

}  // namespace frontend

namespace vs {

// Stream-sharded multi-GPU host layer: stream s lives on GPU s mod G, one host thread + one handle
// (one CUDA context / stream set) per GPU, no collective — only this host-side gather of the per-stream results.
class ShardedTracker {
public:
    ShardedTracker(const vs_tracker_config& base, int n_streams, int n_gpus, int lanes_per_gpu = 4) : n_streams_(n_streams) {
        if (n_gpus < 1 || n_streams < n_gpus) throw std::runtime_error("ShardedTracker: need 1 <= n_gpus <= n_streams");
        shards_.resize(n_gpus);
        for (int g = 0; g < n_gpus; ++g) {
            for (int s = g; s < n_streams; s += n_gpus) shards_[g].streams.push_back(s);
            vs_tracker_config c = base;
            c.device = g;
            c.n_streams = (int)shards_[g].streams.size();
            c.pipeline_lanes = lanes_per_gpu;
            check(vs_tracker_create(&c, nullptr, &shards_[g].handle));
            shards_[g].res.resize(c.n_streams);
            shards_[g].frames.resize(c.n_streams);
            shards_[g].times.resize(c.n_streams);
        }
        results_.resize(n_streams);
    }
    ~ShardedTracker() { for (auto& s : shards_) if (s.handle) vs_tracker_destroy(s.handle); }
    ShardedTracker(const ShardedTracker&) = delete;
    ShardedTracker& operator=(const ShardedTracker&) = delete;

    // frames[s]: H x stride host image of stream s; returns results indexed by global stream id (valid until next call)
    const std::vector<vs_track_result>& track(const uint8_t* const* frames, int stride, const double* time_s) {
        std::vector<std::thread> th;
        std::vector<int> rc(shards_.size(), 0);
        std::vector<std::string> err(shards_.size());
        for (size_t g = 0; g < shards_.size(); ++g) {
            th.emplace_back([&, g] {
                Shard& sh = shards_[g];
                for (size_t k = 0; k < sh.streams.size(); ++k) { sh.frames[k] = frames[sh.streams[k]]; sh.times[k] = time_s[sh.streams[k]]; }
                rc[g] = vs_tracker_track_batch(sh.handle, sh.frames.data(), stride, sh.times.data(), sh.res.data());
                if (rc[g] != VS_OK) err[g] = vs_last_error();          // vs_last_error is thread local
            });
        }
        for (auto& t : th) t.join();
        for (size_t g = 0; g < shards_.size(); ++g) {
            if (rc[g] != VS_OK) throw Error(rc[g], err[g].c_str());
            for (size_t k = 0; k < shards_[g].streams.size(); ++k) results_[shards_[g].streams[k]] = shards_[g].res[k];   // host gather
        }
        return results_;
    }
    int n_gpus() const { return (int)shards_.size(); }
    int n_streams() const { return n_streams_; }

private:
    struct Shard {
        vs_tracker* handle = nullptr;
        std::vector<int> streams;
        std::vector<vs_track_result> res;
        std::vector<const uint8_t*> frames;
        std::vector<double> times;
    };
    std::vector<Shard> shards_;
    std::vector<vs_track_result> results_;
    int n_streams_;
};

}  // namespace vs

#endif  // VS_FEATURE_TRACKER_HPP_
