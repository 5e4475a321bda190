// WorldGather.h — the AoS map -> SoA arrays gather of the reference's loadDataFromWorld (src/adjuster/BundleAdjuster.h:82-98,105-121),
// in canonical indexing, shared by BundleAdjuster.h and IncrementalBundleAdjuster.h.
// Semantics kept: the constant pose is local_frames_.front()->getImage2() (:105); an image that only appears in
// observed_frames_ (the very first one) enters as an all-zero FREE pose (:119, unordered_map::operator[] default-insert);
// intrinsics are shared per Camera (:106-107).  Canonical indexing (the reference's own order is address-hashed, F8d): point id =
// rank of WorldPoint::idx_, pose id = rank of Image::idx_, intrinsics id = first appearance of the Camera by pose id, observation
// id = (point id, position in observed_frames_).
#ifndef B200BA_WORLD_GATHER_H
#define B200BA_WORLD_GATHER_H

#include <algorithm>
#include <cstdint>
#include <unordered_map>
#include <vector>

#include "../world/WorldStructure.h"

struct WorldGather {
    std::vector<Image::Ptr> images_;          // pose id -> image
    std::vector<Camera::Ptr> cameras_;        // intrinsics id -> camera
    std::vector<WorldPoint::Ptr> points_;     // point id -> point
    std::vector<double> pose_, intr_, pt_, obs_uv_;
    std::vector<int32_t> cam_group_, obs_cam_, obs_pt_;
    std::vector<uint8_t> cam_fixed_;

    // The two containers are private to WorldStructure; the friend classes (BundleAdjuster, IncrementalBundleAdjuster) pass them in.
    template <class PointMap, class FrameList>
    void load(const PointMap& world_points, const FrameList& local_frames) {
        // points, by WorldPoint::idx_
        for (auto& kv : world_points) points_.push_back(kv.second);
        std::sort(points_.begin(), points_.end(), [](const WorldPoint::Ptr& a, const WorldPoint::Ptr& b) { return a->idx_ < b->idx_; });
        // poses: image2 of every frame is LOADED (:88-94); every other observed image is default-zero (:119)
        std::unordered_map<Image*, bool> loaded;
        for (const auto& frame : local_frames) {
            auto img = frame->getImage2();
            if (loaded.emplace(img.get(), true).second) images_.push_back(img);
        }
        for (const auto& p : points_)
            for (auto& ob : p->observed_frames_)
                if (loaded.emplace(ob.first.get(), false).second) images_.push_back(ob.first);
        std::sort(images_.begin(), images_.end(), [](const Image::Ptr& a, const Image::Ptr& b) { return a->getIdx() < b->getIdx(); });
        std::unordered_map<Image*, int32_t> pose_id;
        std::unordered_map<Camera*, int32_t> group_id;
        const Image* fixed = local_frames.empty() ? nullptr : local_frames.front()->getImage2().get();
        for (auto& img : images_) {
            pose_id[img.get()] = int32_t(pose_id.size());
            if (loaded[img.get()]) {
                const auto aa = img->getAngleAxisWcMatxCV<double>();
                const auto t = img->getTcw().translation();
                pose_.insert(pose_.end(), {aa(0), aa(1), aa(2), t[0], t[1], t[2]});
            } else {
                pose_.insert(pose_.end(), {0, 0, 0, 0, 0, 0});
            }
            cam_fixed_.push_back(img.get() == fixed ? 1 : 0);
            Camera* cam = img->getCamera().get();
            auto it = group_id.find(cam);
            if (it == group_id.end()) {
                it = group_id.emplace(cam, int32_t(cameras_.size())).first;
                cameras_.push_back(img->getCamera());
                const auto k = img->getCamera()->getIntrinsic();
                intr_.insert(intr_.end(), {k(0), k(1), k(2), k(3)});
            }
            cam_group_.push_back(it->second);
        }
        // points and residual blocks, born sorted by point (:110-121)
        for (size_t p = 0; p < points_.size(); ++p) {
            const auto x = points_[p]->getPosMatx13<double>();
            pt_.insert(pt_.end(), {x(0), x(1), x(2)});
            for (auto& ob : points_[p]->observed_frames_) {
                obs_cam_.push_back(pose_id[ob.first.get()]);
                obs_pt_.push_back(int32_t(p));
                obs_uv_.push_back(ob.second.x);
                obs_uv_.push_back(ob.second.y);
            }
        }
    }

    void clear() {
        images_.clear(); cameras_.clear(); points_.clear();
        pose_.clear(); intr_.clear(); pt_.clear(); obs_uv_.clear();
        cam_group_.clear(); obs_cam_.clear(); obs_pt_.clear(); cam_fixed_.clear();
    }

    // scatter of updateWorld (:143-156): points, poses (incl. the default-inserted first image, as the reference :149-151), intrinsics
    void store() const {
        for (size_t p = 0; p < points_.size(); ++p) points_[p]->setPos(cv::Matx13d(pt_[3 * p], pt_[3 * p + 1], pt_[3 * p + 2]));
        for (size_t c = 0; c < images_.size(); ++c)
            images_[c]->setIntrinsic(cv::Matx23d(pose_[6 * c], pose_[6 * c + 1], pose_[6 * c + 2], pose_[6 * c + 3], pose_[6 * c + 4], pose_[6 * c + 5]));
        for (size_t g = 0; g < cameras_.size(); ++g)
            cameras_[g]->setIntrinsic(cv::Matx14d(intr_[4 * g], intr_[4 * g + 1], intr_[4 * g + 2], intr_[4 * g + 3]));
    }
};

#endif  // B200BA_WORLD_GATHER_H
