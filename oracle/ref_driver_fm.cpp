// TEST INFRASTRUCTURE ONLY — C entry points around the REFERENCE's own compiled FeatureManager
// (oracle/_ref/libvsref_fm.so = /root/reference/src/frontend/feature_manager.cpp, whole and unmodified, compiled where it
// lies against the reference's include/frontend/feature_manager.h and include/utility/config.h, the Eigen / OpenCV stand-ins
// of oracle/ref_shim/ and the two shadow headers of oracle/ref_shim_fm/; see oracle/Makefile, target `ref`).
// Bound with ctypes by oracle/ref_fm_binding.py.  Never linked into, loaded by or shipped with the product.
#include <cstring>
#include <vector>

#include "frontend/feature_manager.h"

extern "C" {
void (*vs_ref_eigvals_cb)(int, const double*, double*, double*) = nullptr;
void (*vs_ref_svd_v_cb)(int, int, const double*, double*) = nullptr;
}

namespace utility {
Config g_config;  // the reference defines it in src/utility/config.cpp (YAML loader; needs cv::FileStorage)
}

using frontend::FeatureManager;

extern "C" {

void vsref_fm_set_svd_callback(void (*cb)(int, int, const double*, double*)) { vs_ref_svd_v_cb = cb; }

void vsref_fm_set_config(double min_parallax, double focal_length, double init_depth) {
    utility::g_config.estimator.min_parallax = min_parallax;
    utility::g_config.camera.focal_length = focal_length;
    utility::g_config.estimator.init_depth = init_depth;
}

int vsref_fm_window_size() { return utility::WINDOW_SIZE; }

void* vsref_fm_create() { return new FeatureManager(); }
void vsref_fm_destroy(void* h) { delete static_cast<FeatureManager*>(h); }
void vsref_fm_clear(void* h) { static_cast<FeatureManager*>(h)->clearState(); }

// feat: n rows of [x, y, z, u, v, vx, vy]; the map orders them by id like common::ImageData
int vsref_fm_add_frame(void* h, int frame_count, int n, const int* ids, const double* feat, int* last_track_num) {
    FeatureManager* fm = static_cast<FeatureManager*>(h);
    common::ImageData image;
    for (int i = 0; i < n; i++) {
        Eigen::Matrix<double, 7, 1> v;
        for (int k = 0; k < 7; k++) v(k) = feat[7 * i + k];
        image[ids[i]] = v;
    }
    const bool key = fm->addFeatureAndCheckParallax(frame_count, image);
    *last_track_num = fm->last_track_num_;
    return key ? 1 : 0;
}

int vsref_fm_bank_size(void* h) { return (int)static_cast<FeatureManager*>(h)->feature_bank_.size(); }
int vsref_fm_feature_count(void* h) { return static_cast<FeatureManager*>(h)->getFeatureCount(); }

// the bank in list order; obs: per feature max_obs rows of 7 doubles [ray xyz, observation uv, velocity]
void vsref_fm_export(void* h, int max_obs, int* ids, int* start_frame, int* n_obs, double* depth, int* solve_flag, double* obs) {
    int e = 0;
    for (auto& it : static_cast<FeatureManager*>(h)->feature_bank_) {
        ids[e] = it.feature_id;
        start_frame[e] = it.start_frame;
        n_obs[e] = (int)it.feature_per_frame.size();
        depth[e] = it.estimated_depth;
        solve_flag[e] = it.solve_flag;
        int k = 0;
        for (auto& f : it.feature_per_frame) {
            if (k >= max_obs) break;
            double* o = obs + ((size_t)e * max_obs + k) * 7;
            o[0] = f.ray_vector.x(); o[1] = f.ray_vector.y(); o[2] = f.ray_vector.z();
            o[3] = f.observation.x(); o[4] = f.observation.y();
            o[5] = f.velocity.x(); o[6] = f.velocity.y();
            k++;
        }
        e++;
    }
}

void vsref_fm_remove_back(void* h) { static_cast<FeatureManager*>(h)->removeBack(); }
void vsref_fm_remove_front(void* h, int frame_count) { static_cast<FeatureManager*>(h)->removeFront(frame_count); }
void vsref_fm_remove_failures(void* h) { static_cast<FeatureManager*>(h)->removeFailures(); }

static Eigen::Matrix3d mat3(const double* r) {  // row-major in
    Eigen::Matrix3d m;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) m(i, j) = r[3 * i + j];
    return m;
}
static Eigen::Vector3d vec3(const double* v) { return Eigen::Vector3d(v[0], v[1], v[2]); }

void vsref_fm_remove_back_shift_depth(void* h, const double* marg_R, const double* marg_P, const double* new_R, const double* new_P) {
    static_cast<FeatureManager*>(h)->removeBackShiftDepth(mat3(marg_R), vec3(marg_P), mat3(new_R), vec3(new_P));
}

// poses: WINDOW_SIZE + 1 rows of [R row-major (9), P (3)]
void vsref_fm_triangulate(void* h, const double* poses, const double* t_ic, const double* r_ic) {
    backend::SlidingWindow w;
    for (int i = 0; i <= utility::WINDOW_SIZE; i++) {
        w[i].R = mat3(poses + 12 * i);
        w[i].P = vec3(poses + 12 * i + 9);
    }
    static_cast<FeatureManager*>(h)->triangulateAcrossAllViews(w, vec3(t_ic), mat3(r_ic));
}

int vsref_fm_get_depth_vector(void* h, double* out) {
    Eigen::VectorXd v = static_cast<FeatureManager*>(h)->getDepthVector();
    for (int i = 0; i < v.rows(); i++) out[i] = v(i);
    return v.rows();
}
static Eigen::VectorXd to_vec(const double* x, int n) {
    Eigen::VectorXd v(n);
    for (int i = 0; i < n; i++) v(i) = x[i];
    return v;
}
void vsref_fm_set_depth(void* h, const double* x, int n) { static_cast<FeatureManager*>(h)->setDepth(to_vec(x, n)); }
void vsref_fm_clear_depth(void* h, const double* x, int n) { static_cast<FeatureManager*>(h)->clearDepth(to_vec(x, n)); }

// pairs: n rows of [a xyz, b xyz]
int vsref_fm_get_corresponding(void* h, int l, int r, int max_n, double* pairs) {
    frontend::Correspondences c = static_cast<FeatureManager*>(h)->getCorresponding(l, r);
    int n = 0;
    for (auto& pr : c) {
        if (n >= max_n) break;
        for (int k = 0; k < 3; k++) {
            pairs[6 * n + k] = pr.first(k);
            pairs[6 * n + 3 + k] = pr.second(k);
        }
        n++;
    }
    return (int)c.size();
}
}
