// BundleAdjuster.h — drop-in host adapter with the reference's call shape
//     BundleAdjuster adj{};  adj(world);                       src/adjuster/BundleAdjuster.h:167-186
// over the C-ABI of the B200 engine (include/ba_engine.h).  It does what the reference's private methods do, in order:
//     loadDataFromWorld  :82-98    gather AoS map -> SoA arrays in canonical indexing
//     bundleAdjustment   :100-141  ba_create / ba_set_problem / ba_solve   (replaces ceres::Solve at :126)
//     updateWorld        :143-156  scatter poses, points and intrinsics back (skipped when the solution is not usable, :180-184)
//     clear              :158-164
// Semantics kept: the constant pose is local_frames_.front()->getImage2() (:105); an image that only appears in
// observed_frames_ (the very first one) enters as an all-zero FREE pose (:119, unordered_map::operator[] default-insert);
// intrinsics are free and shared per Camera (:106-107); one Huber(4) loss on every residual (:109); write-back through
// WorldPoint::setPos, Image::setIntrinsic(Matx23), Camera::setIntrinsic(Matx14).
// Canonical indexing (the reference's own order is address-hashed, F8d): point id = rank of WorldPoint::idx_, pose id = rank
// of Image::idx_, intrinsics id = first appearance of the Camera by pose id, observation id = (point id, position in
// observed_frames_).
// Link with -lba_engine (3dreconstruction_b200/csrc/libba_engine.so).  There is no CPU fallback: without a CUDA device
// ba_create fails, the adapter prints the engine's error and leaves the world untouched (the reference's failure path).
#ifndef B200BA_BUNDLE_ADJUSTER_H
#define B200BA_BUNDLE_ADJUSTER_H

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <iostream>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/ba_engine.h"
#include "../world/WorldStructure.h"

class BundleAdjuster {
    WorldStructure::Ptr structure_;
    ba_options options_{};

    // gathered problem, canonical order
    std::vector<Image::Ptr> images_;          // pose id -> image
    std::vector<Camera::Ptr> cameras_;        // intrinsics id -> camera
    std::vector<WorldPoint::Ptr> points_;     // point id -> point
    std::vector<double> pose_, intr_, pt_, obs_uv_;
    std::vector<int32_t> cam_group_, obs_cam_, obs_pt_;
    std::vector<uint8_t> cam_fixed_;
    ba_summary summary_{};
    std::vector<ba_iter_record> trajectory_;

    void setWorld(const WorldStructure::Ptr& world) { structure_ = world; }

    void loadDataFromWorld() {
        // points, by WorldPoint::idx_
        for (auto& kv : structure_->world_points_) points_.push_back(kv.second);
        std::sort(points_.begin(), points_.end(), [](const WorldPoint::Ptr& a, const WorldPoint::Ptr& b) { return a->idx_ < b->idx_; });
        // poses: image2 of every frame is LOADED (:88-94); every other observed image is default-zero (:119)
        std::unordered_map<Image*, bool> loaded;
        for (const auto& frame : structure_->local_frames_) {
            auto img = frame->getImage2();
            if (loaded.emplace(img.get(), true).second) images_.push_back(img);
        }
        for (const auto& p : points_)
            for (auto& ob : p->observed_frames_)
                if (loaded.emplace(ob.first.get(), false).second) images_.push_back(ob.first);
        std::sort(images_.begin(), images_.end(), [](const Image::Ptr& a, const Image::Ptr& b) { return a->getIdx() < b->getIdx(); });
        std::unordered_map<Image*, int32_t> pose_id;
        std::unordered_map<Camera*, int32_t> group_id;
        const Image* fixed = structure_->local_frames_.empty() ? nullptr : structure_->local_frames_.front()->getImage2().get();
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

    bool bundleAdjustment() {
        ba_problem prob{};
        prob.num_cameras = int32_t(images_.size());
        prob.num_groups = int32_t(cameras_.size());
        prob.num_points = int32_t(points_.size());
        prob.num_observations = int32_t(obs_cam_.size());
        prob.pose = pose_.data(); prob.intr = intr_.data(); prob.cam_group = cam_group_.data(); prob.cam_fixed = cam_fixed_.data();
        prob.pt = pt_.data(); prob.obs_cam = obs_cam_.data(); prob.obs_pt = obs_pt_.data(); prob.obs_uv = obs_uv_.data();

        ba_engine* eng = nullptr;
        int rc = ba_create(&options_, &eng);
        if (rc != BA_OK) {
            std::cout << "Bundle Adjustment failed. (" << ba_last_error(nullptr) << ")" << std::endl;
            return false;
        }
        rc = ba_set_problem(eng, &prob);
        if (rc == BA_OK) rc = ba_solve(eng, &summary_);
        if (rc == BA_OK) {
            int32_t n = 0;
            ba_get_trajectory(eng, nullptr, 0, &n);
            trajectory_.resize(size_t(n));
            if (n) ba_get_trajectory(eng, trajectory_.data(), n, &n);
        }
        if (rc != BA_OK || !summary_.usable) {   // !summary.IsSolutionUsable()  (:128-131)
            std::cout << "Bundle Adjustment failed." << (rc != BA_OK ? std::string(" (") + ba_last_error(eng) + ")" : std::string()) << std::endl;
            ba_destroy(eng);
            return false;
        }
        rc = ba_get_state(eng, pose_.data(), intr_.data(), pt_.data());
        ba_destroy(eng);
        if (rc != BA_OK) { std::cout << "Bundle Adjustment failed." << std::endl; return false; }

        // the reference's statistics block (:134-139)
        std::cout << "Bundle Adjustment statistics (approximated RMSE):\n"
                  << "    #views: " << images_.size() << "\n"
                  << "    #residuals: " << summary_.num_residuals << "\n"
                  << "    Initial RMSE: " << std::sqrt(summary_.initial_cost / summary_.num_residuals) << "\n"
                  << "    Final RMSE: " << std::sqrt(summary_.final_cost / summary_.num_residuals) << "\n"
                  << "    Time (s): " << summary_.solve_seconds << "\n";
        return true;
    }

    void updateWorld() const {
        for (size_t p = 0; p < points_.size(); ++p) points_[p]->setPos(cv::Matx13d(pt_[3 * p], pt_[3 * p + 1], pt_[3 * p + 2]));
        for (size_t c = 0; c < images_.size(); ++c)   // includes the default-inserted first image, as the reference (:149-151)
            images_[c]->setIntrinsic(cv::Matx23d(pose_[6 * c], pose_[6 * c + 1], pose_[6 * c + 2], pose_[6 * c + 3], pose_[6 * c + 4], pose_[6 * c + 5]));
        for (size_t g = 0; g < cameras_.size(); ++g)
            cameras_[g]->setIntrinsic(cv::Matx14d(intr_[4 * g], intr_[4 * g + 1], intr_[4 * g + 2], intr_[4 * g + 3]));
    }

    void clear() {
        structure_ = nullptr;
        images_.clear(); cameras_.clear(); points_.clear();
        pose_.clear(); intr_.clear(); pt_.clear(); obs_uv_.clear();
        cam_group_.clear(); obs_cam_.clear(); obs_pt_.clear(); cam_fixed_.clear();
    }

public:
    [[nodiscard]] BundleAdjuster() {
        // Ceres 2.2 defaults + the reference's settings (:109,:167-174) are the library defaults; the engine solves the
        // reduced camera system by PCG (ITERATIVE_SCHUR with the JACOBI preconditioner the reference already names at :171).
        ba_options_default(&options_);
    }

    // Engine knobs that have no counterpart in the reference's interface (device, PCG tolerance, fixed intrinsics ...).
    ba_options& options() { return options_; }
    // Last solve's summary and per-iteration trajectory (kept across clear(); the reference only prints them).
    const ba_summary& summary() const { return summary_; }
    const std::vector<ba_iter_record>& trajectory() const { return trajectory_; }

    void operator()(WorldStructure::Ptr& world) {
        setWorld(world);
        loadDataFromWorld();
        bool success = bundleAdjustment();
        if (success) updateWorld();
        clear();
    }
};

#endif  // B200BA_BUNDLE_ADJUSTER_H
