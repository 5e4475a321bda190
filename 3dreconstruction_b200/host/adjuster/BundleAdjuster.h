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
#include "WorldGather.h"

class BundleAdjuster : private WorldGather {
    WorldStructure::Ptr structure_;
    ba_options options_{};

    // gathered problem, canonical order (WorldGather.h): images_, cameras_, points_, pose_, intr_, pt_, obs_uv_, cam_group_, obs_cam_, obs_pt_, cam_fixed_
    ba_summary summary_{};
    std::vector<ba_iter_record> trajectory_;

    void setWorld(const WorldStructure::Ptr& world) { structure_ = world; }

    void loadDataFromWorld() { load(structure_->world_points_, structure_->local_frames_); }

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

    void updateWorld() const { store(); }

    void clear() {
        structure_ = nullptr;
        WorldGather::clear();
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
