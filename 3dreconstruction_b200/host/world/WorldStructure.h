// WorldStructure.h — the map container the adjuster mutates in place (mirror of src/world/WorldStructure.h:27-156 without
// PCL I/O and the viewer: graphics are unavailable on this box).  `friend class BundleAdjuster` as the reference (:38).
#ifndef B200BA_WORLD_STRUCTURE_H
#define B200BA_WORLD_STRUCTURE_H
#include <memory>
#include <unordered_map>
#include <vector>

#include "../component/Image.h"
#include "../frame/LocalFrame.h"
#include "WorldPoint.h"

class WorldStructure {
public:
    using Ptr = std::shared_ptr<WorldStructure>;

private:
    WorldPoint::Idx cur_idx_ = 0;
    std::vector<LocalFrame::Ptr> local_frames_;
    std::unordered_map<Image::Idx, Image::Ptr> images_{};
    std::unordered_map<WorldPoint::Idx, WorldPoint::Ptr> world_points_{};

    friend class BundleAdjuster;
    friend class IncrementalBundleAdjuster;   // the one line a maintainer adds for the persistent adjuster

public:
    WorldStructure() = default;
    auto addImage(Image::Ptr& image) { images_[image->getIdx()] = image; }
    auto addPoint(Eigen::Vector3d world_pos) -> std::size_t {   // reference :47-52 (colour, descriptor dropped)
        auto point_idx = cur_idx_++;
        world_points_[point_idx] = std::make_shared<WorldPoint>(point_idx, world_pos);
        return point_idx;
    }
    auto addLocalFrame(LocalFrame::Ptr local_frame) { local_frames_.emplace_back(std::move(local_frame)); }
    auto getLocalFrames() -> const std::vector<LocalFrame::Ptr>& { return local_frames_; }
    auto getPointFromIdx(const WorldPoint::Idx idx) const { return world_points_.at(idx); }
    auto getWorldPoints() const -> const std::unordered_map<WorldPoint::Idx, WorldPoint::Ptr>& { return world_points_; }
    auto getImages() const -> const std::unordered_map<Image::Idx, Image::Ptr>& { return images_; }
};
#endif
