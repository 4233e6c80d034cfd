// vs_feature_manager.hpp — header-only C++ host layer over the device feature bank (vs_bank_* of include/vs_tracker.h).
//
// `frontend::FeatureManager` below has the method names, argument meaning and public members of the class the
// reference's estimator uses (/root/reference/include/frontend/feature_manager.h:70-92, methods in
// /root/reference/src/frontend/feature_manager.cpp), for ONE camera stream; the features live on the GPU.  Many streams
// share launches through the C ABI directly (one vs_bank handle, per-stream arguments).
//
// Types.  Without Eigen the class speaks plain arrays: vs::Vec3 (3 doubles), vs::Mat3 (9 doubles, row-major),
// vs::ImageData (std::map<int, std::array<double, 7>>: id -> [x, y, z, u, v, vx, vy], common/image_frame.h:14) and
// vs::Poses (WINDOW_SIZE + 1 pairs R, P).  Define VS_WITH_EIGEN before including this header (the reference's build has
// Eigen) and the reference's own signatures exist too: common::ImageData with Eigen::Matrix<double, 7, 1>,
// Eigen::Matrix3d / Vector3d / VectorXd, and any window type with operator[] whose elements have .R and .P
// (backend::SlidingWindow).
//
// `feature_bank_` (read directly by estimator.cpp:363,396) is a host MIRROR of the device list, refreshed after every
// call that changes the bank while `mirror_feature_bank` is true (default; one device-to-host copy per call), or on demand
// with refreshFeatureBank().  There is no CPU path: the constructor throws vs::BankError without a CUDA device.
#ifndef VS_FEATURE_MANAGER_HPP_
#define VS_FEATURE_MANAGER_HPP_

#include <array>
#include <list>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "vs_tracker.h"
#ifdef VS_WITH_REFERENCE_CONFIG
#include "utility/config.h"   // utility::g_config of the reference
#endif

namespace vs {

using Vec3 = std::array<double, 3>;
using Mat3 = std::array<double, 9>;                       // row-major
using ImageData = std::map<int, std::array<double, 7>>;
using Poses = std::array<std::pair<Mat3, Vec3>, VS_BANK_MAX_OBS>;

class BankError : public std::runtime_error {
public:
    BankError(int code_, const std::string& what) : std::runtime_error(what), code(code_) {}
    int code;
};

// what the estimator reads of a feature (feature_manager.h:25-68)
struct FeaturePerFrame {
    Vec3 ray_vector;
    std::array<double, 2> observation;
    std::array<double, 2> velocity;
};
struct FeaturePerId {
    int feature_id = 0;
    int start_frame = 0;
    std::vector<FeaturePerFrame> feature_per_frame;
    int used_num = 0;
    double estimated_depth = -1.0;
    int solve_flag = 0;   // 0 haven't solve yet; 1 solve succ; 2 solve fail
    int endFrame() const { return start_frame + (int)feature_per_frame.size() - 1; }
};

}  // namespace vs

namespace frontend {

using Correspondence = std::pair<vs::Vec3, vs::Vec3>;
using Correspondences = std::vector<Correspondence>;

class FeatureManager {
public:
    // the three values the reference reads from utility::g_config (estimator.min_parallax, camera.focal_length,
    // estimator.init_depth; include/utility/config.h:19,63-64) are members here; with VS_WITH_REFERENCE_CONFIG they are
    // re-read from g_config on every call like the reference does
    double min_parallax = 10.0, focal_length = 460.0, init_depth = 5.0;
    bool mirror_feature_bank = true;
    std::list<vs::FeaturePerId> feature_bank_;
    int last_track_num_ = 0;

    explicit FeatureManager(int capacity = 2048, int device = 0) : capacity_(capacity) {
        check(vs_bank_create(device, 1, capacity, &bank_));
    }
    ~FeatureManager() { vs_bank_destroy(bank_); }
    FeatureManager(const FeatureManager&) = delete;
    FeatureManager& operator=(const FeatureManager&) = delete;

    void clearState() {                                                    // feature_manager.cpp:11-13
        check(vs_bank_clear(bank_, 0));
        changed();
    }

    int getFeatureCount() {                                                // :15-25
        int32_t n = 0;
        check(vs_bank_get_depth_vector(bank_, 0, 0, nullptr, &n));
        return n;
    }

    bool addFeatureAndCheckParallax(int frame_count, const vs::ImageData& image) {   // :27-73
        read_config();
        const int32_t n = (int32_t)image.size();
        ids_.resize(n > 0 ? n : 1);
        rec_.resize(7 * (size_t)(n > 0 ? n : 1));
        int k = 0;
        for (const auto& kv : image) {
            ids_[k] = kv.first;
            for (int q = 0; q < 7; ++q) rec_[7 * (size_t)k + q] = kv.second[q];
            ++k;
        }
        const int32_t fc = frame_count;
        int32_t key = 0, tracked = 0;
        check(vs_bank_add_frame(bank_, &fc, &n, ids_.data(), rec_.data(), n, min_parallax, focal_length, &key, &tracked,
                                &last_parallax_sum, &last_parallax_num));
        last_track_num_ = tracked;
        changed();
        return key != 0;
    }

    Correspondences getCorresponding(int frame_count_l, int frame_count_r) {          // :75-92
        int32_t n = 0;
        check(vs_bank_get_corresponding(bank_, 0, frame_count_l, frame_count_r, 0, nullptr, &n));
        std::vector<double> p(6 * (size_t)(n > 0 ? n : 1));
        check(vs_bank_get_corresponding(bank_, 0, frame_count_l, frame_count_r, n, p.data(), &n));
        Correspondences c((size_t)n);
        for (int i = 0; i < n; ++i)
            for (int q = 0; q < 3; ++q) {
                c[i].first[q] = p[6 * (size_t)i + q];
                c[i].second[q] = p[6 * (size_t)i + 3 + q];
            }
        return c;
    }

    void setDepth(const std::vector<double>& x) {                          // :94-110
        check(vs_bank_set_depth(bank_, 0, x.data(), (int32_t)x.size(), 0));
        changed();
    }
    void removeFailures() { slide(4, 0, nullptr); }                        // :112-118
    void clearDepth(const std::vector<double>& x) {                        // :120-128
        check(vs_bank_set_depth(bank_, 0, x.data(), (int32_t)x.size(), 1));
        changed();
    }
    std::vector<double> getDepthVector() {                                 // :130-140
        int32_t n = 0;
        std::vector<double> v((size_t)capacity_);
        check(vs_bank_get_depth_vector(bank_, 0, capacity_, v.data(), &n));
        v.resize((size_t)n);
        return v;
    }

    void triangulateAcrossAllViews(const vs::Poses& window, const vs::Vec3& t_ic, const vs::Mat3& r_ic) {   // :142-194
        read_config();
        double poses[VS_BANK_MAX_OBS * 12];
        for (int i = 0; i < VS_BANK_MAX_OBS; ++i) {
            for (int q = 0; q < 9; ++q) poses[12 * i + q] = window[i].first[q];
            for (int q = 0; q < 3; ++q) poses[12 * i + 9 + q] = window[i].second[q];
        }
        check(vs_bank_triangulate(bank_, poses, t_ic.data(), r_ic.data(), init_depth, nullptr));
        changed();
    }

    void removeBackShiftDepth(const vs::Mat3& marg_R, const vs::Vec3& marg_P, const vs::Mat3& new_R, const vs::Vec3& new_P) {   // :208-235
        read_config();
        double rp[24];
        for (int q = 0; q < 9; ++q) { rp[q] = marg_R[q]; rp[12 + q] = new_R[q]; }
        for (int q = 0; q < 3; ++q) { rp[9 + q] = marg_P[q]; rp[21 + q] = new_P[q]; }
        slide(3, 0, rp);
    }
    void removeBack() { slide(1, 0, nullptr); }                            // :237-249
    void removeFront(int frame_count) { slide(2, frame_count, nullptr); }  // :251-266
    // removeOutlier (:196-206) is `assert(false)` in the reference and has no caller: not provided

    // refresh the host mirror `feature_bank_` from the device
    void refreshFeatureBank() {
        int32_t n = 0;
        const size_t C = (size_t)capacity_;
        m_ids_.resize(C); m_start_.resize(C); m_nobs_.resize(C); m_flag_.resize(C); m_depth_.resize(C);
        m_obs_.resize(C * VS_BANK_MAX_OBS * 7);
        check(vs_bank_export(bank_, 0, capacity_, m_ids_.data(), m_start_.data(), m_nobs_.data(), m_depth_.data(),
                             m_flag_.data(), m_obs_.data(), &n));
        feature_bank_.clear();
        for (int i = 0; i < n; ++i) {
            vs::FeaturePerId f;
            f.feature_id = m_ids_[i];
            f.start_frame = m_start_[i];
            f.estimated_depth = m_depth_[i];
            f.solve_flag = m_flag_[i];
            f.used_num = m_nobs_[i];
            f.feature_per_frame.resize((size_t)m_nobs_[i]);
            for (int k = 0; k < m_nobs_[i]; ++k) {
                const double* o = m_obs_.data() + ((size_t)i * VS_BANK_MAX_OBS + k) * 7;
                f.feature_per_frame[k].ray_vector = {o[0], o[1], o[2]};
                f.feature_per_frame[k].observation = {o[3], o[4]};
                f.feature_per_frame[k].velocity = {o[5], o[6]};
            }
            feature_bank_.push_back(std::move(f));
        }
    }

    // the parallax sum / count the last key-frame decision was taken on (diagnostics; 0 / 0 when the early-outs fired)
    double last_parallax_sum = 0.0;
    int32_t last_parallax_num = 0;
    struct vs_bank* handle() { return bank_; }

#ifdef VS_WITH_EIGEN
    // ---- the reference's own signatures (feature_manager.h:76-88) ----
    template <class ImageMap>   // common::ImageData = std::map<int, Eigen::Matrix<double, 7, 1>>
    bool addFeatureAndCheckParallax(int frame_count, const ImageMap& image) {
        vs::ImageData im;
        for (const auto& kv : image) {
            std::array<double, 7> r;
            for (int q = 0; q < 7; ++q) r[q] = kv.second(q);
            im[kv.first] = r;
        }
        return addFeatureAndCheckParallax(frame_count, im);
    }
    template <class VecX>
    void setDepth(const VecX& x) { setDepth(to_std(x)); }
    template <class VecX>
    void clearDepth(const VecX& x) { clearDepth(to_std(x)); }
    template <class Window, class V3, class M3>   // backend::SlidingWindow, Eigen::Vector3d, Eigen::Matrix3d
    void triangulateAcrossAllViews(const Window& w, const V3& t_ic, const M3& r_ic) {
        vs::Poses p;
        for (int i = 0; i < VS_BANK_MAX_OBS; ++i) p[i] = std::make_pair(mat3(w[i].R), vec3(w[i].P));
        triangulateAcrossAllViews(p, vec3(t_ic), mat3(r_ic));
    }
    template <class M3, class V3>
    void removeBackShiftDepth(const M3& marg_R, const V3& marg_P, const M3& new_R, const V3& new_P) {
        removeBackShiftDepth(mat3(marg_R), vec3(marg_P), mat3(new_R), vec3(new_P));
    }
#endif

private:
    struct vs_bank* bank_ = nullptr;
    int capacity_;
    std::vector<int32_t> ids_, m_ids_, m_start_, m_nobs_, m_flag_;
    std::vector<double> rec_, m_depth_, m_obs_;

    static void check(int rc) {
        if (rc != VS_OK) throw vs::BankError(rc, std::string("vs_bank: ") + vs_last_error());
    }
    void changed() {
        if (mirror_feature_bank) refreshFeatureBank();
    }
    void slide(int32_t op, int32_t frame_count, const double* rp) {
        read_config();
        check(vs_bank_slide(bank_, &op, &frame_count, rp, init_depth));
        changed();
    }
    void read_config() {
#ifdef VS_WITH_REFERENCE_CONFIG
        min_parallax = utility::g_config.estimator.min_parallax;
        focal_length = utility::g_config.camera.focal_length;
        init_depth = utility::g_config.estimator.init_depth;
#endif
    }
#ifdef VS_WITH_EIGEN
    template <class M3>
    static vs::Mat3 mat3(const M3& m) {
        vs::Mat3 o;
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) o[3 * i + j] = m(i, j);
        return o;
    }
    template <class V3>
    static vs::Vec3 vec3(const V3& v) { return vs::Vec3{v(0), v(1), v(2)}; }
    template <class VecX>
    static std::vector<double> to_std(const VecX& x) {
        std::vector<double> v((size_t)x.rows());
        for (int i = 0; i < (int)x.rows(); ++i) v[i] = x(i);
        return v;
    }
#endif
};

}  // namespace frontend

#endif  // VS_FEATURE_MANAGER_HPP_
