// oracle/residuals.h — CPU restatement of the reference's residual models and robust kernels.
// TEST INFRASTRUCTURE ONLY.  The two cost functors below are checked against the reference's OWN functors compiled from
// /root/reference (oracle/ref_build.sh -> oracle/_ref/libref_cost.so; tests/test_reference_functors.py: equal to 1e-13, vectors committed
// under tests/golden/).  Still unpinned: the third-party pieces (AngleAxisRotatePoint, HuberLoss, CauchyLoss of Ceres Solver 2.2,
// un-vendored: CMakeLists.txt:32), restated from their published algorithms and cross-checked against cv2.Rodrigues /
// cv2.projectPoints / finite differences in tests/test_oracle_kat.py.
#ifndef BA_ORACLE_RESIDUALS_H
#define BA_ORACLE_RESIDUALS_H
#include <cfloat>
#include <cmath>
#include <limits>

#include "jet.h"

namespace oracle {

// ceres::AngleAxisRotatePoint (call sites: src/adjuster/BundleAdjuster.h:45, SnavelyReprojectionError.h:23).
// Away from zero: Rodrigues, result = p cos(th) + (w x p) sin(th) + w (w.p)(1 - cos(th)), w = aa/th.
// At zero: first-order R = I + [aa]x so that the Jet derivative stays finite (d/daa = -[p]x).
// taylor_rule 0: branch on th^2 <= DBL_EPSILON (Ceres <= 2.1 form); 1: branch on hypot(aa) == 0 (2.2 form).
template <typename T>
inline void angle_axis_rotate_point(const T aa[3], const T p[3], T out[3], int taylor_rule) {
    bool rodrigues;
    T theta;
    if (taylor_rule == 0) {
        const T theta2 = aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2];
        rodrigues = value_of(theta2) > std::numeric_limits<double>::epsilon();
        if (rodrigues) theta = sqrt(theta2);
    } else {
        rodrigues = std::hypot(value_of(aa[0]), value_of(aa[1]), value_of(aa[2])) != 0.0;
        if (rodrigues) theta = hypot3(aa[0], aa[1], aa[2]);
    }
    if (rodrigues) {
        const T c = cos(theta), s = sin(theta), inv = T(1.0) / theta;
        const T w[3] = {aa[0] * inv, aa[1] * inv, aa[2] * inv};
        const T wxp[3] = {w[1] * p[2] - w[2] * p[1], w[2] * p[0] - w[0] * p[2], w[0] * p[1] - w[1] * p[0]};
        const T k = (w[0] * p[0] + w[1] * p[1] + w[2] * p[2]) * (T(1.0) - c);
        out[0] = p[0] * c + wxp[0] * s + w[0] * k;
        out[1] = p[1] * c + wxp[1] * s + w[1] * k;
        out[2] = p[2] * c + wxp[2] * s + w[2] * k;
    } else {
        out[0] = p[0] + (aa[1] * p[2] - aa[2] * p[1]);
        out[1] = p[1] + (aa[2] * p[0] - aa[0] * p[2]);
        out[2] = p[2] + (aa[0] * p[1] - aa[1] * p[0]);
    }
}

// ReprojectCost::operator() — src/adjuster/BundleAdjuster.h:38-68.
// Blocks: intrinsic(4: fx,fy,cx,cy) :55-58, extrinsic(6: angle-axis, t) :41-42, point(3).
// P = R(aa) X + t (:45-50); +z forward, plain perspective divide (:52-53); u = fx x + cx, v = fy y + cy (:61-62);
// residual = prediction - observation (:64-65).  No cheirality test; always "succeeds".
struct PinholeCost {
    static constexpr int kIntr = 4;
    double ox, oy;
    int taylor_rule;
    template <typename T>
    void operator()(const T* intr, const T* ext, const T* X, T* res) const {
        T P[3];
        angle_axis_rotate_point(ext, X, P, taylor_rule);
        P[0] += ext[3]; P[1] += ext[4]; P[2] += ext[5];
        const T xn = P[0] / P[2];
        const T yn = P[1] / P[2];
        res[0] = intr[0] * xn + intr[2] - T(ox);
        res[1] = intr[1] * yn + intr[3] - T(oy);
    }
};

// SnavelyReprojectionError::operator() — src/adjuster/SnavelyReprojectionError.h:16-48 (never instantiated by the
// reference; the BAL model).  Blocks: camera(3: f,l1,l2) :17, image(6) :18, point(3) :19.
// -z camera (:30-31), radial distortion 1 + r2 (l1 + l2 r2) (:34-37), prediction f d (xp,yp) (:40-42).
struct SnavelyCost {
    static constexpr int kIntr = 3;
    double ox, oy;
    int taylor_rule;
    template <typename T>
    void operator()(const T* camera, const T* image, const T* X, T* res) const {
        T P[3];
        angle_axis_rotate_point(image, X, P, taylor_rule);
        P[0] += image[3]; P[1] += image[4]; P[2] += image[5];
        const T xp = -P[0] / P[2];
        const T yp = -P[1] / P[2];
        const T r2 = xp * xp + yp * yp;
        const T d = 1.0 + r2 * (camera[1] + camera[2] * r2);
        res[0] = camera[0] * d * xp - T(ox);
        res[1] = camera[0] * d * yp - T(oy);
    }
};

// Robust kernels on s = ||r||^2: rho[0] = rho(s), rho[1] = rho'(s), rho[2] = rho''(s).
// ceres::HuberLoss(a) (BundleAdjuster.h:109, a = 4) and ceres::CauchyLoss(a); SURVEY a6 / KAT-5.
inline void loss_evaluate(int kind, double a, double s, double rho[3]) {
    const double b = a * a;
    if (kind == 1) {  // Huber
        if (s > b) {
            const double r = std::sqrt(s);
            rho[0] = 2.0 * a * r - b;
            rho[1] = std::max(std::numeric_limits<double>::min(), a / r);
            rho[2] = -rho[1] / (2.0 * s);
        } else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
    } else if (kind == 2) {  // Cauchy
        const double c = 1.0 / b, sum = 1.0 + s * c, inv = 1.0 / sum;
        rho[0] = b * std::log(sum);
        rho[1] = std::max(std::numeric_limits<double>::min(), inv);
        rho[2] = -c * (inv * inv);
    } else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
}

}  // namespace oracle
#endif
