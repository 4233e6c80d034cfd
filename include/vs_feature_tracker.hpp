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
#include <cmath>
#include <cstring>
#include <fstream>
#include <condition_variable>
#include <map>
#include <memory>
#include <mutex>
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
    std::string fisheye_mask;             // g_config.feature_tracker.fisheye_mask: mask image file (feature_tracker.cpp:42-46)
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
        return (size_t)img.step;
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
        c.fisheye_mask = g.feature_tracker.fisheye_mask;
#endif
        return c;
    }
    void reset_handle() { if (handle_) { vs_tracker_destroy(handle_); handle_ = nullptr; } }
    // feature_tracker.cpp:42-46: the mask file is read once, lazily, the first time a frame is tracked with `fisheye` set;
    // an unreadable or wrongly sized file falls back to the generated circle (there: `fisheye_mask.empty()`)
    bool load_mask_image(const vs::FrontendConfig& c, int cols, int rows, std::vector<uint8_t>* out) {
        out->clear();
        if (!c.fisheye || c.fisheye_mask.empty()) return false;
#ifdef VS_WITH_OPENCV
        cv::Mat m = cv::imread(c.fisheye_mask, cv::IMREAD_GRAYSCALE);
        if (m.empty() || m.cols != cols || m.rows != rows) return false;
        out->resize((size_t)cols * rows);
        for (int y = 0; y < rows; ++y) std::memcpy(out->data() + (size_t)y * cols, m.ptr<uint8_t>(y), (size_t)cols);
        return true;
#else
        int32_t w = 0, h = 0;
        out->resize((size_t)cols * rows);
        if (vs_image_read_gray(c.fisheye_mask.c_str(), out->data(), cols, cols, rows, &w, &h) != VS_OK || w != cols || h != rows) {
            out->clear();
            return false;
        }
        return true;
#endif
    }
    void ensure_handle(int cols, int rows) {
        const vs::FrontendConfig c = current();
        // the reference takes inBorder limits, mask size and the virtual-pinhole centre from g_config.camera.row/col and
        // everything else from the image; its callers always hand over rowRange(0, camera.row) of a camera.col-wide image
        // (measurement_processor.cpp:191), and OpenCV throws in calcOpticalFlowPyrLK when the size changes mid-run
        if (c.col != cols || c.row != rows)
            throw std::runtime_error("FeatureTracker: image is " + std::to_string(cols) + "x" + std::to_string(rows) +
                                     " but camera.col x camera.row is " + std::to_string(c.col) + "x" + std::to_string(c.row));
        if (handle_) {
            if (cols != handle_cols_ || rows != handle_rows_)
                throw std::runtime_error("FeatureTracker: image size changed from " + std::to_string(handle_cols_) + "x" +
                                         std::to_string(handle_rows_) + " to " + std::to_string(cols) + "x" + std::to_string(rows));
            return;
        }
        if (!have_camera_) throw std::runtime_error("FeatureTracker: camera intrinsics not set");
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
        std::vector<uint8_t> mask;
        if (load_mask_image(c, cols, rows, &mask)) { tc.fisheye_mask_image = mask.data(); tc.fisheye_mask_stride = cols; mask_loaded_ = true; }
        vs::check(vs_tracker_create(&tc, nullptr, &handle_));
        handle_cols_ = cols; handle_rows_ = rows;
        applied_ = c;
#ifdef VS_WITH_OPENCV
        if (c.ransac_mode == VS_RANSAC_CALLBACK) vs::check(vs_tracker_set_reject_callback(handle_, &FeatureTracker::reject_cv, this));
#endif
    }
    // Every call re-reads the configuration like the reference does (g_config is read inside detectAndTrack / setMask /
    // rejectWithFundamentalMatrix): the VIOEngine mutators (vio_engine.cpp:354-389) and direct writes to equalize,
    // fisheye and camera.focal_length all take effect on the next frame without losing the tracking state.
    void apply_runtime_params() {
        const vs::FrontendConfig c = current();
        if (c.max_cnt != applied_.max_cnt || c.min_dist != applied_.min_dist || c.f_threshold != applied_.f_threshold ||
            c.lk_window_size != applied_.lk_window_size || c.lk_pyramid_levels != applied_.lk_pyramid_levels ||
            c.lk_iterations != applied_.lk_iterations || c.lk_eps != applied_.lk_eps ||
            c.f_threshold_edge_factor != applied_.f_threshold_edge_factor) {
            vs::check(vs_tracker_set_params(handle_, c.max_cnt, c.min_dist, c.f_threshold, c.lk_window_size, c.lk_pyramid_levels,
                                            c.lk_iterations, c.lk_eps, c.f_threshold_edge_factor));
            // vs_tracker_set_params keeps the guards of VIOEngine::setTrackingParams (values <= 0, even windows and
            // levels > 5 leave the field unchanged): say so when a value written straight into the config was not taken
            vs_tracker_config eff;
            vs::check(vs_tracker_get_config(handle_, &eff));
            auto warn = [](const char* name, double want, double have) {
                if (want != have) std::fprintf(stderr, "[vs] %s = %g is not supported at run time, the tracker keeps %g\n", name, want, have);
            };
            warn("max_cnt", c.max_cnt, eff.max_cnt); warn("min_dist", c.min_dist, eff.min_dist);
            warn("f_threshold", c.f_threshold, eff.f_threshold); warn("lk_window_size", c.lk_window_size, eff.lk_window_size);
            warn("lk_pyramid_levels", c.lk_pyramid_levels, eff.lk_pyramid_levels);
            warn("lk_iterations", c.lk_iterations, eff.lk_iterations); warn("lk_eps", c.lk_eps, eff.lk_eps);
            warn("f_threshold_edge_factor", c.f_threshold_edge_factor, eff.f_threshold_edge_factor);
        }
        if (c.equalize != applied_.equalize || c.fisheye != applied_.fisheye || c.focal_length != applied_.focal_length) {
            std::vector<uint8_t> mask;
            const bool first_fisheye = c.fisheye && !applied_.fisheye && !mask_loaded_ && load_mask_image(c, handle_cols_, handle_rows_, &mask);
            vs::check(vs_tracker_set_frontend_flags(handle_, c.equalize ? 1 : 0, c.fisheye ? 1 : 0, c.focal_length,
                                                    first_fisheye ? mask.data() : nullptr, handle_cols_));
            if (first_fisheye) mask_loaded_ = true;
        }
        applied_ = c;
    }
#ifdef VS_WITH_OPENCV
    // rejectWithFundamentalMatrix, feature_tracker.cpp:226-299: the lift to the virtual pinhole was done on the GPU,
    // the RANSAC itself is the reference's own OpenCV call, edge recovery as in :236-285
    static int reject_cv(void* user, int32_t, int32_t n, const float* a, const float* b, const float* px, uint8_t* status) {
        auto* self = static_cast<FeatureTracker*>(user);
        const vs::FrontendConfig c = self->current();
        std::vector<cv::Point2f> A(n), B(n);
        for (int i = 0; i < n; ++i) { A[i] = cv::Point2f(a[2 * i], a[2 * i + 1]); B[i] = cv::Point2f(b[2 * i], b[2 * i + 1]); }
        std::vector<uchar> st;
        cv::Mat F = cv::findFundamentalMat(A, B, cv::FM_RANSAC, c.f_threshold, 0.99, st);
        for (int i = 0; i < n; ++i) status[i] = i < (int)st.size() ? st[i] : 0;
        const double edge_factor = c.f_threshold_edge_factor;
        if (edge_factor > 0.0 && !F.empty() && F.rows == 3 && F.cols == 3) {          // feature_tracker.cpp:236-285
            const double cx = c.col / 2.0, cy = c.row / 2.0;
            const double r_max = std::sqrt(cx * cx + cy * cy);
            const double r_max_inv = (r_max > 1e-6) ? 1.0 / r_max : 0.0;
            const double* f = F.ptr<double>();
            for (int i = 0; i < n; ++i) {
                if (status[i]) continue;
                const double dx = px[2 * i] - cx, dy = px[2 * i + 1] - cy;
                const double r_ratio = std::sqrt(dx * dx + dy * dy) * r_max_inv;
                if (r_ratio < 0.3) continue;
                const double u1 = A[i].x, v1 = A[i].y, u2 = B[i].x, v2 = B[i].y;
                const double la = f[0] * u1 + f[1] * v1 + f[2];
                const double lb = f[3] * u1 + f[4] * v1 + f[5];
                const double lc = f[6] * u1 + f[7] * v1 + f[8];
                const double norm_ab = std::sqrt(la * la + lb * lb);
                if (norm_ab < 1e-12) continue;
                const double dist = std::abs(la * u2 + lb * v2 + lc) / norm_ab;
                if (dist < c.f_threshold * (1.0 + edge_factor * r_ratio * r_ratio)) status[i] = 1;
            }
        }
        return 0;
    }
#endif
    vs_tracker* handle_ = nullptr;
    vs_camera camera_{};
    bool have_camera_ = false;
    bool mask_loaded_ = false;
    int handle_cols_ = 0, handle_rows_ = 0;
    vs::FrontendConfig applied_{};
    std::vector<int> real_ids_;
};

}  // namespace frontend

namespace vs {

// ---- the callers' side of the path -------------------------------------------------------------
// What MeasurementProcessor::extractImageFeatures (src/utility/measurement_processor.cpp:178-231) returns for one frame:
// rays (id, x, y, 1) plus the five channels id, u, v, vx, vy of every feature with track_cnt > 1 (ImageFeatureMsg).
struct ImageFeatureMsg {
    double timestamp = 0.0;
    std::string frame_id;
    int points_count = 0;
    int channels_count = 0;
    struct Point3D { int id; double x, y, z; };
    std::vector<Point3D> ray_vectors;
    std::vector<double> channel_data[5];       // id, u, v, velocity_x, velocity_y
};

// the publish loop shared by both callers (measurement_processor.cpp:200-231, vio_engine.cpp:219-239)
inline void publish(const frontend::FeatureTracker& ft, ImageFeatureMsg* msg) {
    for (size_t j = 0; j < ft.ids.size(); ++j) {
        if (ft.track_cnt[j] <= 1) continue;
        const int id = ft.ids[j];
        msg->ray_vectors.push_back({id, ft.cur_undistorted_pts[j].x, ft.cur_undistorted_pts[j].y, 1.0});
        msg->channel_data[0].push_back(id);
        msg->channel_data[1].push_back(ft.cur_pts[j].x);
        msg->channel_data[2].push_back(ft.cur_pts[j].y);
        msg->channel_data[3].push_back(ft.pts_velocity[j].x);
        msg->channel_data[4].push_back(ft.pts_velocity[j].y);
    }
    msg->points_count = (int)msg->ray_vectors.size();
    msg->channels_count = 5;
}

// MeasurementProcessor::extractImageFeatures: read the file as grey (cv::imread(path, IMREAD_GRAYSCALE) with OpenCV, the
// library's own PNG / PGM decoder without), track rowRange(0, camera.row), run the updateID loop, publish.  An unreadable
// file returns the message with only timestamp and frame_id set, like the reference ("Cannot load image").
inline ImageFeatureMsg extractImageFeatures(frontend::FeatureTracker& ft, const std::string& full_path, const std::string& filename,
                                            double timestamp, int camera_row, int camera_col) {
    ImageFeatureMsg msg;
    msg.timestamp = timestamp;
    msg.frame_id = filename;
#ifdef VS_WITH_OPENCV
    cv::Mat show_img = cv::imread(full_path, cv::IMREAD_GRAYSCALE);
    if (show_img.empty()) { std::fprintf(stderr, "Cannot load image: %s\n", full_path.c_str()); return msg; }
    ft.detectAndTrack(show_img.rowRange(0, camera_row), timestamp);
    (void)camera_col;
#else
    std::vector<uint8_t> buf((size_t)camera_row * camera_col);
    int32_t w = 0, h = 0;
    if (vs_image_read_gray(full_path.c_str(), buf.data(), camera_col, camera_col, camera_row, &w, &h) != VS_OK || w != camera_col || h < camera_row) {
        std::fprintf(stderr, "Cannot load image: %s (%s)\n", full_path.c_str(), vs_last_error());
        return msg;
    }
    ft.detectAndTrack(frontend::FeatureTracker::Image(buf.data(), camera_col, camera_row), timestamp);
#endif
    for (unsigned int i = 0;; i++)
        if (!ft.updateID(i)) break;
    publish(ft, &msg);
    return msg;
}

// The front-end half of VIOEngine::processFrame (src/vio_engine.cpp:180, 210-239): wrap the caller's grey buffer without
// copying, track, run the updateID loop and build common::ImageData (id -> [x, y, 1, u, v, vx, vy]).
inline std::map<int, std::vector<double>> processFrameFrontend(frontend::FeatureTracker& ft, const uint8_t* gray_image, int width,
                                                               int height, double image_timestamp) {
#ifdef VS_WITH_OPENCV
    cv::Mat img(height, width, CV_8UC1, const_cast<uint8_t*>(gray_image));
#else
    frontend::FeatureTracker::Image img(gray_image, width, height);
#endif
    ft.detectAndTrack(img, image_timestamp);
    for (unsigned int i = 0;; i++)
        if (!ft.updateID(i)) break;
    return ft.imageData();
}

// Stream-sharded multi-GPU host layer: stream s lives on GPU s mod G, one handle (one CUDA context /
// stream set) and ONE PERSISTENT WORKER THREAD per GPU, no collective — only the host-side gather of the per-stream
// results.  submit() / wait() map onto vs_tracker_submit_batch / vs_tracker_wait of every shard (two batches may be in
// flight, so the upload of batch k+1 overlaps the kernels of batch k on every GPU); track() = submit() + wait().
class ShardedTracker {
public:
    ShardedTracker(const vs_tracker_config& base, int n_streams, int n_gpus, int lanes_per_gpu = 4) : n_streams_(n_streams) {
        if (n_gpus < 1 || n_streams < n_gpus) throw std::runtime_error("ShardedTracker: need 1 <= n_gpus <= n_streams");
        for (int g = 0; g < n_gpus; ++g) shards_.emplace_back(new Shard());
        for (int g = 0; g < n_gpus; ++g) {
            Shard& sh = *shards_[g];
            for (int s = g; s < n_streams; s += n_gpus) sh.streams.push_back(s);
            vs_tracker_config c = base;
            c.device = g;
            c.n_streams = (int)sh.streams.size();
            c.pipeline_lanes = lanes_per_gpu;
            const int rc = vs_tracker_create(&c, nullptr, &sh.handle);
            if (rc != VS_OK) { const Error e(rc, vs_last_error()); shutdown(); throw e; }
            sh.res.resize(c.n_streams);
            for (int k = 0; k < 2; ++k) { sh.frames[k].resize(c.n_streams); sh.times[k].resize(c.n_streams); }
        }
        for (auto& sp : shards_) { Shard* sh = sp.get(); sh->worker = std::thread([this, sh] { worker_loop(*sh); }); }
        results_.resize(n_streams);
    }
    ~ShardedTracker() { shutdown(); }
    ShardedTracker(const ShardedTracker&) = delete;
    ShardedTracker& operator=(const ShardedTracker&) = delete;

    // frames[s]: H x stride host image of stream s; must stay valid until the matching wait() returns
    void submit(const uint8_t* const* frames, int stride, const double* time_s) {
        if (submitted_ - waited_ >= 2) throw std::runtime_error("ShardedTracker: two batches already in flight, call wait()");
        const int slot = (int)(submitted_ & 1);
        for (auto& sp : shards_) {
            Shard& sh = *sp;
            for (size_t k = 0; k < sh.streams.size(); ++k) { sh.frames[slot][k] = frames[sh.streams[k]]; sh.times[slot][k] = time_s[sh.streams[k]]; }
            post(sh, slot, stride, kSubmit);
        }
        join_all();
        ++submitted_;
    }
    // results of the OLDEST submitted batch, indexed by global stream id (valid until the second wait() after this one)
    const std::vector<vs_track_result>& wait() {
        if (submitted_ == waited_) throw std::runtime_error("ShardedTracker: nothing in flight");
        for (auto& sp : shards_) post(*sp, 0, 0, kWait);
        join_all();
        ++waited_;
        for (auto& sp : shards_)
            for (size_t k = 0; k < sp->streams.size(); ++k) results_[sp->streams[k]] = sp->res[k];       // the host-side gather
        return results_;
    }
    const std::vector<vs_track_result>& track(const uint8_t* const* frames, int stride, const double* time_s) {
        submit(frames, stride, time_s);
        return wait();
    }
    int n_gpus() const { return (int)shards_.size(); }
    int n_streams() const { return n_streams_; }

private:
    enum Cmd { kIdle = 0, kSubmit, kWait, kQuit };
    struct Shard {
        vs_tracker* handle = nullptr;
        std::vector<int> streams;
        std::vector<vs_track_result> res;
        std::vector<const uint8_t*> frames[2];
        std::vector<double> times[2];
        std::thread worker;
        std::mutex mu;
        std::condition_variable cv;
        int cmd = kIdle, slot = 0, stride = 0;
        bool done = true;
        int rc = VS_OK;
        std::string err;
    };
    void post(Shard& sh, int slot, int stride, int cmd) {
        std::lock_guard<std::mutex> lk(sh.mu);
        sh.cmd = cmd; sh.slot = slot; sh.stride = stride; sh.done = false;
        sh.cv.notify_all();
    }
    void join_all() {
        int rc = VS_OK;
        std::string err;
        for (auto& sp : shards_) {
            Shard& sh = *sp;
            std::unique_lock<std::mutex> lk(sh.mu);
            sh.cv.wait(lk, [&] { return sh.done; });
            if (sh.rc != VS_OK && rc == VS_OK) { rc = sh.rc; err = sh.err; }
        }
        if (rc != VS_OK) throw Error(rc, err.c_str());
    }
    void worker_loop(Shard& sh) {
        for (;;) {
            int cmd, slot, stride;
            {
                std::unique_lock<std::mutex> lk(sh.mu);
                sh.cv.wait(lk, [&] { return sh.cmd != kIdle; });
                cmd = sh.cmd; slot = sh.slot; stride = sh.stride;
                sh.cmd = kIdle;
            }
            if (cmd == kQuit) return;
            int rc = cmd == kSubmit ? vs_tracker_submit_batch(sh.handle, sh.frames[slot].data(), stride, sh.times[slot].data())
                                    : vs_tracker_wait(sh.handle, sh.res.data());
            std::lock_guard<std::mutex> lk(sh.mu);
            sh.rc = rc;
            if (rc != VS_OK) sh.err = vs_last_error();          // vs_last_error is thread local
            sh.done = true;
            sh.cv.notify_all();
        }
    }
    void shutdown() {
        for (auto& sp : shards_) {
            Shard& sh = *sp;
            if (sh.worker.joinable()) { post(sh, 0, 0, kQuit); sh.worker.join(); }
            if (sh.handle) { vs_tracker_destroy(sh.handle); sh.handle = nullptr; }
        }
    }
    std::vector<std::unique_ptr<Shard>> shards_;      // Shard holds a mutex: not movable
    std::vector<vs_track_result> results_;
    int n_streams_;
    int64_t submitted_ = 0, waited_ = 0;
};

}  // namespace vs

#endif  // VS_FEATURE_TRACKER_HPP_
