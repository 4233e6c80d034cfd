// TEST INFRASTRUCTURE ONLY — stand-in for <ceres/rotation.h> (Ceres is not installed).  The reference's camera
// headers name ceres::QuaternionRotatePoint inside spaceToPlane templates that the tracker path never instantiates.
#ifndef VS_REF_SHIM_CERES_ROTATION
#define VS_REF_SHIM_CERES_ROTATION
namespace ceres {
template <class T>
void QuaternionRotatePoint(const T q[4], const T pt[3], T result[3]);
}
#endif
