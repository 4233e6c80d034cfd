// TEST INFRASTRUCTURE ONLY — stand-in for the reference's include/common/image_frame.h (feature_manager.cpp needs only the
// ImageData alias, image_frame.h:14; the class behind it drags in the IMU factor headers).  Our code, not reference code.
#ifndef VS_REF_SHIM_IMAGE_FRAME_H
#define VS_REF_SHIM_IMAGE_FRAME_H
#include <Eigen/Dense>
#include <map>

namespace common {
using ImageData = std::map<int, Eigen::Matrix<double, 7, 1>>;
}
#endif
