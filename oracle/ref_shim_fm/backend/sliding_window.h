// TEST INFRASTRUCTURE ONLY — stand-in for the reference's include/backend/sliding_window.h, used only when
// /root/reference/src/frontend/feature_manager.cpp is compiled into oracle/_ref/libvsref_fm.so.  The reference's header
// pulls in the IMU pre-integration and Ceres factor headers (backend/factor/*.h), none of which
// FeatureManager::triangulateAcrossAllViews touches: it reads sliding_window[i].P and .R and nothing else
// (feature_manager.cpp:158-168).  Our code, not reference code.
#ifndef VS_REF_SHIM_SLIDING_WINDOW_H
#define VS_REF_SHIM_SLIDING_WINDOW_H
#include <Eigen/Dense>
#include <array>

#include "utility/config.h"

// the reference's backend/factor/integration_base.h:10-11 (reached through the real sliding_window.h) opens these two
// namespaces at global scope, and feature_manager.cpp relies on that for WINDOW_SIZE and g_config
using namespace Eigen;
using namespace utility;

namespace common {
struct Frame {
    Eigen::Matrix3d R;
    Eigen::Vector3d P;
};
}  // namespace common

namespace backend {
class SlidingWindow {
public:
    common::Frame& operator[](int index) { return sliding_window[index]; }
    const common::Frame& operator[](int index) const { return sliding_window[index]; }

private:
    std::array<common::Frame, WINDOW_SIZE + 1> sliding_window;
};
}  // namespace backend
#endif
