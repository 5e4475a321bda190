// lite_types.h — dependency-free stand-ins for the handful of OpenCV / Eigen / Sophus types that the reference's map
// classes expose to the adjuster (src/world, src/component, src/frame).  Only what the bundle-adjustment path touches is
// here: cv::Point2d, cv::Matx<T,1,3>/<T,1,4>/<T,2,3>, Eigen::Vector3d and Sophus::SE3d (rotation matrix + translation).
// In the reference tree these are the real libraries' types; the adapter (host/adjuster/BundleAdjuster.h) uses only the
// members declared here, so swapping the real headers back in needs no change to it.
#ifndef B200BA_LITE_TYPES_H
#define B200BA_LITE_TYPES_H
#include <array>
#include <cmath>
#include <cstddef>

namespace cv {
struct Point2d { double x = 0, y = 0; Point2d() = default; Point2d(double x_, double y_) : x(x_), y(y_) {} };
struct Point2f { float x = 0, y = 0; Point2f() = default; Point2f(float x_, float y_) : x(x_), y(y_) {}
                 operator Point2d() const { return Point2d(x, y); } };   // cv::KeyPoint::pt -> cv::Point2d (SequentialActuator.h:41,66-67)
template <class T, int M, int N>
struct Matx {
    T val[M * N] = {};
    Matx() = default;
    template <class... A, class = std::enable_if_t<sizeof...(A) == size_t(M * N)>>
    Matx(A... a) : val{T(a)...} {}
    T& operator()(int i, int j) { return val[i * N + j]; }
    const T& operator()(int i, int j) const { return val[i * N + j]; }
    T& operator()(int i) { return val[i]; }
    const T& operator()(int i) const { return val[i]; }
};
using Matx13d = Matx<double, 1, 3>;
using Matx14d = Matx<double, 1, 4>;
using Matx23d = Matx<double, 2, 3>;
}  // namespace cv

namespace Eigen {
struct Vector3d {
    double v[3] = {0, 0, 0};
    Vector3d() = default;
    Vector3d(double x, double y, double z) : v{x, y, z} {}
    double& operator()(int i) { return v[i]; }
    const double& operator()(int i) const { return v[i]; }
    double& operator()(int i, int) { return v[i]; }
    const double& operator()(int i, int) const { return v[i]; }
    double& operator[](int i) { return v[i]; }
    const double& operator[](int i) const { return v[i]; }
};
}  // namespace Eigen

namespace Sophus {
// World-to-camera rigid transform: row-major rotation matrix and translation.
class SE3d {
    double R_[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    Eigen::Vector3d t_;
public:
    SE3d() = default;
    SE3d(const double* R, const Eigen::Vector3d& t) : t_(t) { for (int i = 0; i < 9; ++i) R_[i] = R[i]; }
    const double* rotationMatrix() const { return R_; }
    void setRotationMatrix(const double* R) { for (int i = 0; i < 9; ++i) R_[i] = R[i]; }
    Eigen::Vector3d& translation() { return t_; }
    const Eigen::Vector3d& translation() const { return t_; }
};
}  // namespace Sophus

namespace b200ba {
// R = exp([aa]x) (Rodrigues), row-major.
inline void angle_axis_to_rotation(const double aa[3], double R[9]) {
    const double th2 = aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2];
    if (th2 > 0.0) {
        const double th = std::sqrt(th2), c = std::cos(th), s = std::sin(th), k = 1.0 - c;
        const double x = aa[0] / th, y = aa[1] / th, z = aa[2] / th;
        R[0] = c + k * x * x;     R[1] = k * x * y - s * z; R[2] = k * x * z + s * y;
        R[3] = k * y * x + s * z; R[4] = c + k * y * y;     R[5] = k * y * z - s * x;
        R[6] = k * z * x - s * y; R[7] = k * z * y + s * x; R[8] = c + k * z * z;
    } else {
        R[0] = R[4] = R[8] = 1.0; R[1] = R[2] = R[3] = R[5] = R[6] = R[7] = 0.0;
    }
}

// angle * axis of a rotation matrix, through the unit quaternion as Eigen::AngleAxisd(Matrix3d) does (Image.h:87-93);
// angle in [0, pi].
inline void rotation_to_angle_axis(const double R[9], double aa[3]) {
    double q[4];   // w x y z
    const double tr = R[0] + R[4] + R[8];
    if (tr > 0.0) {
        double t = std::sqrt(tr + 1.0);
        q[0] = 0.5 * t; t = 0.5 / t;
        q[1] = (R[7] - R[5]) * t; q[2] = (R[2] - R[6]) * t; q[3] = (R[3] - R[1]) * t;
    } else {
        int i = 0;
        if (R[4] > R[0]) i = 1;
        if (R[8] > R[4 * i]) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        double t = std::sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
        q[1 + i] = 0.5 * t; t = 0.5 / t;
        q[0] = (R[3 * k + j] - R[3 * j + k]) * t;
        q[1 + j] = (R[3 * j + i] + R[3 * i + j]) * t;
        q[1 + k] = (R[3 * k + i] + R[3 * i + k]) * t;
    }
    if (q[0] < 0.0) for (double& v : q) v = -v;
    const double n = std::sqrt(q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    if (n == 0.0) { aa[0] = aa[1] = aa[2] = 0.0; return; }
    const double angle = 2.0 * std::atan2(n, q[0]);
    aa[0] = angle * q[1] / n; aa[1] = angle * q[2] / n; aa[2] = angle * q[3] / n;
}

// Rz(a) * Ry(b) * Rx(c): what Image::setIntrinsic builds from the optimised triple (src/component/Image.h:130-141).
inline void euler_zyx_to_rotation(double a, double b, double c, double R[9]) {
    const double ca = std::cos(a), sa = std::sin(a), cb = std::cos(b), sb = std::sin(b), cc = std::cos(c), sc = std::sin(c);
    R[0] = ca * cb; R[1] = ca * sb * sc - sa * cc; R[2] = ca * sb * cc + sa * sc;
    R[3] = sa * cb; R[4] = sa * sb * sc + ca * cc; R[5] = sa * sb * cc - ca * sc;
    R[6] = -sb;     R[7] = cb * sc;                R[8] = cb * cc;
}
}  // namespace b200ba
#endif
