// Image.h — a view: global index, camera, world-to-camera pose (mirror of src/component/Image.h:16-173, BA-relevant part;
// pixels, keypoints and descriptors dropped).
#ifndef B200BA_IMAGE_H
#define B200BA_IMAGE_H
#include <memory>
#include <string>

#include "../common/lite_types.h"
#include "Camera.h"

class Image {
public:
    using Idx = std::size_t;
    using Ptr = std::shared_ptr<Image>;
    // How setIntrinsic(Matx23) interprets the optimised rotation triple.  The reference builds Rz(a0) Ry(a1) Rx(a2) from an
    // ANGLE-AXIS triple (Image.h:130-141) — a first-order bug.  Default here: the correct exponential map.
    enum class WriteBack { CorrectAngleAxis, ReferenceEulerBug };
    static inline WriteBack write_back = WriteBack::CorrectAngleAxis;

private:
    static inline Idx cur_idx_ = 0;   // global counter, as the reference (:22,:36)
    Idx idx_;
    std::string name_;
    Camera::Ptr camera_;
    Sophus::SE3d Tcw;                 // transform from world to camera

public:
    [[nodiscard]] explicit Image(std::string image_path, Camera::Ptr camera) : idx_{cur_idx_++}, name_{std::move(image_path)}, camera_{std::move(camera)} {}

    static auto getTotalImages() { return cur_idx_; }
    auto getName() const { return name_; }
    auto getCamera() { return camera_; }
    auto getIdx() const { return idx_; }

    template <typename T>
    cv::Matx<T, 1, 3> getAngleAxisWcMatxCV() const {   // reference :87-93 (Eigen::AngleAxisd(R): angle * axis)
        double aa[3];
        b200ba::rotation_to_angle_axis(Tcw.rotationMatrix(), aa);
        return cv::Matx<T, 1, 3>(aa[0], aa[1], aa[2]);
    }
    [[nodiscard]] Sophus::SE3d getTcw() const { return Tcw; }
    void setTcw(const Sophus::SE3d& tcw) { Tcw = tcw; }

    template <typename T>
    void setIntrinsic(cv::Matx<T, 2, 3> angleAxisAndTranslation) {   // reference :130-141 (the pose write-back, despite its name)
        double R[9];
        if (write_back == WriteBack::ReferenceEulerBug) {
            b200ba::euler_zyx_to_rotation(angleAxisAndTranslation(0, 0), angleAxisAndTranslation(0, 1), angleAxisAndTranslation(0, 2), R);
        } else {
            const double aa[3] = {double(angleAxisAndTranslation(0, 0)), double(angleAxisAndTranslation(0, 1)), double(angleAxisAndTranslation(0, 2))};
            b200ba::angle_axis_to_rotation(aa, R);
        }
        Tcw.setRotationMatrix(R);
        Tcw.translation() = Eigen::Vector3d(angleAxisAndTranslation(1, 0), angleAxisAndTranslation(1, 1), angleAxisAndTranslation(1, 2));
    }
    [[nodiscard]] auto getIntrinsic() const {   // reference :143-151
        const auto aa = getAngleAxisWcMatxCV<double>();
        const auto& t = Tcw.translation();
        return cv::Matx<double, 2, 3>(aa(0, 0), aa(0, 1), aa(0, 2), t(0), t(1), t(2));
    }
};
#endif
