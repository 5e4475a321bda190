// oracle/shim/opencv2/core/types.hpp — cv::Point2d as the reference's ReprojectCost stores it (BundleAdjuster.h:34-36). TEST INFRASTRUCTURE ONLY.
#ifndef BA_SHIM_OPENCV_TYPES_HPP
#define BA_SHIM_OPENCV_TYPES_HPP
namespace cv {
template <typename T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
};
using Point2d = Point_<double>;
}  // namespace cv
#endif
