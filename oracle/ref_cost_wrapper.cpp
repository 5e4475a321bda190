// oracle/ref_cost_wrapper.cpp — C entry points around the REFERENCE's own residual functors, compiled from the sources where they
// lie under /root/reference (never copied into this repository):
//   * SnavelyReprojectionError   — /root/reference/src/adjuster/SnavelyReprojectionError.h, included unmodified;
//   * BundleAdjuster::ReprojectCost — a private nested class of src/adjuster/BundleAdjuster.h:33-69, whose enclosing header needs
//     OpenCV, PCL, Sophus, Eigen and spdlog; oracle/ref_build.sh cuts exactly that class out of the reference file into
//     oracle/_ref/reproject_cost.inc (git-ignored, rebuilt from the reference every time) and it is included below.
// The functors are differentiated with the same forward-mode Jets the oracle uses (Ceres' AutoDiffCostFunction does the same with its
// own Jet); ceres::AngleAxisRotatePoint comes from oracle/shim (Ceres is not in this image).
// TEST INFRASTRUCTURE ONLY: used by tests/golden/make_reference_vectors.py and tests/test_reference_functors.py to pin the oracle's
// (and, through committed vectors, the engine's) residuals and Jacobians to the reference's source.
#include <ceres/ceres.h>
#include <ceres/rotation.h>
#include <opencv2/core/types.hpp>

#include "adjuster/SnavelyReprojectionError.h"   // -I/root/reference/src

namespace ref {
#include "_ref/reproject_cost.inc"                // class ReprojectCost { ... };  (BundleAdjuster.h:33-69)
}

namespace {
template <int NI, class Functor>
void differentiate(const Functor& f, const double* intr, const double* ext, const double* X, double res[2], double* J) {
    constexpr int N = NI + 9;
    using JetT = oracle::Jet<N>;
    JetT a[NI], b[6], c[3], r[2];
    for (int k = 0; k < NI; ++k) a[k] = JetT(intr[k], k);
    for (int k = 0; k < 6; ++k) b[k] = JetT(ext[k], NI + k);
    for (int k = 0; k < 3; ++k) c[k] = JetT(X[k], NI + 6 + k);
    f(a, b, c, r);
    for (int q = 0; q < 2; ++q) { res[q] = r[q].a; for (int k = 0; k < N; ++k) J[q * N + k] = r[q].v[k]; }
}
}  // namespace

extern "C" {
// ReprojectCost(observation)(intrinsic[4], extrinsic[6], pos3d[3]) -> residuals[2]; J is 2 x 13 row-major, columns (fx fy cx cy | aa t | X)
int ref_reproject_cost(const double* intr, const double* ext, const double* X, const double* obs, int taylor_rule, double* res, double* J) {
    ceres::shim_taylor_rule() = taylor_rule;
    const ref::ReprojectCost f(cv::Point2d(obs[0], obs[1]));
    differentiate<4>(f, intr, ext, X, res, J);
    double plain[2];
    // the plain double instantiation agrees with the Jet one (to rounding: a Jet divides by multiplying with the reciprocal, as Ceres' does)
    auto close = [](double a, double b) { return std::fabs(a - b) <= 1e-9 * (1.0 + std::fabs(b)); };
    return f(intr, ext, X, plain) && close(plain[0], res[0]) && close(plain[1], res[1]) ? 0 : 1;
}
// SnavelyReprojectionError(x, y)(camera[3], image[6], point[3]) -> residuals[2]; J is 2 x 12 row-major, columns (f l1 l2 | aa t | X)
int ref_snavely_cost(const double* cam, const double* image, const double* X, const double* obs, int taylor_rule, double* res, double* J) {
    ceres::shim_taylor_rule() = taylor_rule;
    const SnavelyReprojectionError f(obs[0], obs[1]);
    differentiate<3>(f, cam, image, X, res, J);
    ceres::CostFunction* cf = SnavelyReprojectionError::Create(obs[0], obs[1]);           // the reference's factory compiles and runs
    delete cf;
    return 0;
}
}
