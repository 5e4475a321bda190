// Camera.h — the intrinsics holder the adjuster reads and writes (mirror of src/component/Camera.h:11-121, BA-relevant part).
#ifndef B200BA_CAMERA_H
#define B200BA_CAMERA_H
#include <memory>

#include "../common/lite_types.h"

class Camera {
public:
    using Ptr = std::shared_ptr<Camera>;

private:
    double fx, fy, cx, cy;

public:
    Camera(const double fx, const double fy, const double cx, const double cy) : fx(fx), fy(fy), cx(cx), cy(cy) {}

    template <typename T>
    void setIntrinsic(cv::Matx<T, 1, 4> intrinsic) {   // reference Camera.h:29-35
        fx = intrinsic(0); fy = intrinsic(1); cx = intrinsic(2); cy = intrinsic(3);
    }
    [[nodiscard]] auto getIntrinsic() const { return cv::Matx14d(fx, fy, cx, cy); }   // reference Camera.h:94-96
    [[nodiscard]] double get_fx() const { return fx; }
    [[nodiscard]] double get_fy() const { return fy; }
    [[nodiscard]] double get_cx() const { return cx; }
    [[nodiscard]] double get_cy() const { return cy; }
};
#endif
