// TEST INFRASTRUCTURE ONLY — C entry points around the REFERENCE's own compiled tracker (oracle/_ref/libvsref.so).
//
// Linked with (see oracle/Makefile, target `ref`):
//   * /root/reference/src/frontend/feature_tracker.cpp                     — compiled where it lies, unmodified, whole;
//   * oracle/_ref/gen/ref_*.cc                                              — the liftProjective-path and YAML-loading functions
//     of the reference's camera models and CameraFactory, copied verbatim at build time by oracle/ref_extract.py
//     (git-ignored output, one unit per reference file);
//   * the reference's own headers (include/frontend/feature_tracker.h, include/common/camera_models/*.h,
//     include/utility/config.h) resolved against the Eigen / OpenCV / Ceres stand-ins in oracle/ref_shim/.
// This file holds only what the link needs beyond that: the `utility::g_config` global, aborting stubs for the
// virtuals of the camera classes that are NOT on the tracker's path (calibration, projection, YAML writing, undistortion
// maps — they need the real OpenCV/Eigen SDKs), and the `vsref_*` functions Python binds with ctypes
// (oracle/ref_binding.py).  Never linked into, loaded by or shipped with the product.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "frontend/feature_tracker.h"
#include "common/camera_models/CataCamera.h"
#include "common/camera_models/ScaramuzzaCamera.h"

extern "C" {
vs_ref_callbacks vs_ref_cb = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
void (*vs_ref_eigvals_cb)(int, const double*, double*, double*) = nullptr;
}

namespace utility {
Config g_config;  // the reference defines it in src/utility/config.cpp (YAML loader; needs cv::FileStorage)
}

[[noreturn]] static void off_path(const char* what) {
    std::fprintf(stderr, "oracle/_ref: %s is not on the tracker path and is not built here\n", what);
    std::abort();
}

namespace common {
namespace camera_models {

// ---- not on the path: stubs so the vtables link ------------------------------------------------------------------
void Camera::estimateExtrinsics(const std::vector<cv::Point3f>&, const std::vector<cv::Point2f>&, cv::Mat&, cv::Mat&) const {
    off_path("Camera::estimateExtrinsics");
}

#define VSREF_STUBS(CLS)                                                                                                \
    void CLS::Parameters::writeToYamlFile(const std::string&) const { off_path(#CLS "::Parameters::writeToYamlFile"); } \
    void CLS::estimateIntrinsics(const cv::Size&, const std::vector<std::vector<cv::Point3f>>&,                         \
                                 const std::vector<std::vector<cv::Point2f>>&) {                                        \
        off_path(#CLS "::estimateIntrinsics");                                                                          \
    }                                                                                                                   \
    void CLS::spaceToPlane(const Eigen::Vector3d&, Eigen::Vector2d&) const { off_path(#CLS "::spaceToPlane"); }         \
    void CLS::undistToPlane(const Eigen::Vector2d&, Eigen::Vector2d&) const { off_path(#CLS "::undistToPlane"); }       \
    cv::Mat CLS::initUndistortRectifyMap(cv::Mat&, cv::Mat&, float, float, cv::Size, float, float, cv::Mat) const {     \
        off_path(#CLS "::initUndistortRectifyMap");                                                                     \
    }                                                                                                                   \
    void CLS::writeParametersToYamlFile(const std::string&) const { off_path(#CLS "::writeParametersToYamlFile"); }     \
    std::string CLS::parametersToString(void) const { off_path(#CLS "::parametersToString"); }

VSREF_STUBS(PinholeCamera)
VSREF_STUBS(CataCamera)
VSREF_STUBS(EquidistantCamera)
VSREF_STUBS(OCAMCamera)

// CameraFactory::instance / generateCameraFromYamlFile and the cameras' readFromYamlFile come from the reference (generated
// units); the static member's definition line is not a function body, so it is repeated here (CameraFactory.cc:27)
std::shared_ptr<CameraFactory> CameraFactory::m_instance;

}  // namespace camera_models
}  // namespace common

using common::camera_models::CataCamera;
using common::camera_models::OCAMCamera;
using frontend::FeatureTracker;

struct vsref_config {
    double row, col, focal_length;
    int max_cnt, min_dist;
    double f_threshold;
    int equalize, fisheye;
    int lk_window_size, lk_pyramid_levels, lk_iterations;
    double lk_eps, f_threshold_edge_factor;
};

extern "C" {

void vsref_set_callbacks(const vs_ref_callbacks* cb) {
    vs_ref_cb = *cb;
    vs_ref_eigvals_cb = cb->eigvals;
}

// utility::g_config is process-global in the reference (include/utility/config.h:118); the binding sets it before
// every call so several trackers can coexist in one test process
void vsref_set_config(const vsref_config* c, const char* fisheye_mask_path) {
    auto& g = utility::g_config;
    g.camera.row = c->row;
    g.camera.col = c->col;
    g.camera.focal_length = c->focal_length;
    auto& f = g.feature_tracker;
    f.max_cnt = c->max_cnt;
    f.min_dist = c->min_dist;
    f.f_threshold = c->f_threshold;
    f.equalize = c->equalize;
    f.fisheye = c->fisheye;
    f.lk_window_size = c->lk_window_size;
    f.lk_pyramid_levels = c->lk_pyramid_levels;
    f.lk_iterations = c->lk_iterations;
    f.lk_eps = c->lk_eps;
    f.f_threshold_edge_factor = c->f_threshold_edge_factor;
    f.fisheye_mask = fisheye_mask_path ? fisheye_mask_path : "";
}

void vsref_set_n_id(int v) { FeatureTracker::n_id = v; }
int vsref_get_n_id(void) { return FeatureTracker::n_id; }

void* vsref_create(void) { return new FeatureTracker(); }
void vsref_destroy(void* h) { delete (FeatureTracker*)h; }

// FeatureTracker::setIntrinsicParameter (feature_tracker.cpp:322-336): PINHOLE or (anything else) KANNALA_BRANDT
void vsref_set_intrinsic(void* h, int model_type, int w, int hgt, double fx, double fy, double cx, double cy, double k2,
                         double k3, double k4, double k5) {
    ((FeatureTracker*)h)->setIntrinsicParameter(model_type, w, hgt, fx, fy, cx, cy, k2, k3, k4, k5);
}
// the YAML route builds these through CameraFactory; here the same camera classes are constructed from their
// Parameters (the reference's constructors) and handed to the public m_camera member
void vsref_set_camera_mei(void* h, int w, int hgt, double xi, double k1, double k2, double p1, double p2, double g1,
                          double g2, double u0, double v0) {
    CataCamera::Parameters p("camera", w, hgt, xi, k1, k2, p1, p2, g1, g2, u0, v0);
    ((FeatureTracker*)h)->m_camera = std::make_shared<CataCamera>(p);
}
void vsref_set_camera_ocam(void* h, int w, int hgt, double C, double D, double E, double cx, double cy, const double* poly5) {
    OCAMCamera::Parameters p;
    p.imageWidth() = w;
    p.imageHeight() = hgt;
    p.C() = C;
    p.D() = D;
    p.E() = E;
    p.center_x() = cx;
    p.center_y() = cy;
    for (int i = 0; i < SCARAMUZZA_POLY_SIZE; i++) p.poly(i) = poly5[i];
    ((FeatureTracker*)h)->m_camera = std::make_shared<OCAMCamera>(p);
}

// FeatureTracker::readIntrinsicParameter (feature_tracker.cpp:318-320) -> CameraFactory::generateCameraFromYamlFile
// (CameraFactory.cc:87-150) -> <Model>::Parameters::readFromYamlFile, all reference code; cv::FileStorage is answered
// by cv2.FileStorage.  Returns the model type, or -1 when the factory returned a null camera.
int vsref_read_intrinsic(void* h, const char* yaml_path) {
    FeatureTracker* ft = (FeatureTracker*)h;
    ft->readIntrinsicParameter(yaml_path);
    return ft->m_camera ? (int)ft->m_camera->modelType() : -1;
}
// the camera's parameter vector (Camera::writeParameters) and image size; returns the parameter count
int vsref_camera_parameters(void* h, double* out, int cap, int* width, int* height) {
    FeatureTracker* ft = (FeatureTracker*)h;
    if (!ft->m_camera) return -1;
    std::vector<double> p;
    ft->m_camera->writeParameters(p);
    for (int i = 0; i < (int)p.size() && i < cap; ++i) out[i] = p[i];
    *width = ft->m_camera->imageWidth();
    *height = ft->m_camera->imageHeight();
    return (int)p.size();
}

// FeatureTracker::detectAndTrack on a caller-owned grey image (vio_engine.cpp:180 wraps the raw pointer the same way)
void vsref_detect_and_track(void* h, const unsigned char* gray, int rows, int cols, long step, double t) {
    cv::Mat img(rows, cols, CV_8UC1, (void*)gray, (size_t)step);
    ((FeatureTracker*)h)->detectAndTrack(img, t);
}

// the callers' id loop (measurement_processor.cpp:193-198, vio_engine.cpp:213-216)
int vsref_update_ids(void* h) {
    FeatureTracker* ft = (FeatureTracker*)h;
    unsigned int i = 0;
    for (;; i++)
        if (!ft->updateID(i)) break;
    return (int)i;
}

int vsref_count(void* h) { return (int)((FeatureTracker*)h)->ids.size(); }

// copies of the five public vectors the callers read (any pointer may be NULL)
void vsref_get(void* h, int* ids, int* track_cnt, float* cur_pts, float* un_pts, float* vel) {
    FeatureTracker* ft = (FeatureTracker*)h;
    const size_t n = ft->ids.size();
    for (size_t i = 0; i < n; i++) {
        if (ids) ids[i] = ft->ids[i];
        if (track_cnt) track_cnt[i] = ft->track_cnt[i];
        if (cur_pts) {
            cur_pts[2 * i] = ft->cur_pts[i].x;
            cur_pts[2 * i + 1] = ft->cur_pts[i].y;
        }
        if (un_pts) {
            un_pts[2 * i] = ft->cur_undistorted_pts[i].x;
            un_pts[2 * i + 1] = ft->cur_undistorted_pts[i].y;
        }
        if (vel) {
            vel[2 * i] = ft->pts_velocity[i].x;
            vel[2 * i + 1] = ft->pts_velocity[i].y;
        }
    }
}

// the mask after setMask (rows*cols bytes, packed); returns 0 when the tracker holds none yet
int vsref_get_mask(void* h, unsigned char* out) {
    FeatureTracker* ft = (FeatureTracker*)h;
    if (ft->mask.empty()) return 0;
    for (int y = 0; y < ft->mask.rows; y++)
        std::memcpy(out + (size_t)y * ft->mask.cols, ft->mask.data + (size_t)y * ft->mask.step, ft->mask.cols);
    return 1;
}

int vsref_n_new(void* h) { return (int)((FeatureTracker*)h)->n_pts.size(); }

// Camera::liftProjective of the tracker's camera over n pixels
void vsref_lift_projective(void* h, int n, const double* uv, double* xyz) {
    FeatureTracker* ft = (FeatureTracker*)h;
    for (int i = 0; i < n; i++) {
        Eigen::Vector3d P;
        ft->m_camera->liftProjective(Eigen::Vector2d(uv[2 * i], uv[2 * i + 1]), P);
        xyz[3 * i] = P.x();
        xyz[3 * i + 1] = P.y();
        xyz[3 * i + 2] = P.z();
    }
}

int vsref_in_border(float x, float y) { return frontend::inBorder(cv::Point2f(x, y)) ? 1 : 0; }

}  // extern "C"
