// WorldPoint.h — a map point with its observation list (mirror of src/world/WorldPoint.h:13-41; colour/descriptor dropped).
#ifndef B200BA_WORLDPOINT_H
#define B200BA_WORLDPOINT_H
#include <cstddef>
#include <memory>
#include <utility>
#include <vector>

#include "../common/lite_types.h"

class Image;

struct WorldPoint {
    using Ptr = std::shared_ptr<WorldPoint>;
    using Idx = std::size_t;
    static constexpr Idx InvalidIdx = static_cast<Idx>(-1);

    Idx idx_;
    Eigen::Vector3d world_pos_;
    std::vector<std::pair<std::shared_ptr<Image>, cv::Point2d>> observed_frames_;   // reference :24

    WorldPoint(Idx idx, const Eigen::Vector3d& pos) : idx_(idx), world_pos_(pos) {}

    template <typename T>
    cv::Matx<T, 1, 3> getPosMatx13() const { return cv::Matx<T, 1, 3>(world_pos_(0, 0), world_pos_(1, 0), world_pos_(2, 0)); }   // :30-33
    template <typename T>
    void setPos(cv::Matx<T, 1, 3> p) { world_pos_(0) = p(0); world_pos_(1) = p(1); world_pos_(2) = p(2); }                           // :35-40
};
#endif
