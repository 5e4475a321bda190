// oracle/shim/ceres/rotation.h — stand-in for <ceres/rotation.h> so that the REFERENCE's own cost functors
// (/root/reference/src/adjuster/SnavelyReprojectionError.h, and the ReprojectCost class of BundleAdjuster.h) compile here unmodified.
// TEST INFRASTRUCTURE ONLY (oracle/ref_build.sh).  Ceres Solver itself is not in this image and not vendored by the reference
// (CMakeLists.txt:32); ceres::AngleAxisRotatePoint below is the oracle's restatement of its published algorithm (oracle/residuals.h).
#ifndef BA_SHIM_CERES_ROTATION_H
#define BA_SHIM_CERES_ROTATION_H
#include "../../residuals.h"

namespace ceres {
using oracle::Jet;
// which small-angle rule the shim applies (0: theta^2 <= DBL_EPSILON, 1: hypot == 0); set by the wrapper before a call
inline int& shim_taylor_rule() { static int rule = 0; return rule; }

template <typename T>
inline void AngleAxisRotatePoint(const T angle_axis[3], const T pt[3], T result[3]) {
    oracle::angle_axis_rotate_point(angle_axis, pt, result, shim_taylor_rule());
}
}  // namespace ceres

// the reference's functors call sqrt / sin / cos unqualified on T in some Ceres versions; make the Jet overloads visible
using oracle::sqrt; using oracle::sin; using oracle::cos;
#endif
