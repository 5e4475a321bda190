// IncrementalBundleAdjuster.h — the same call shape as BundleAdjuster (`adj(world);`), for the way the reference actually uses bundle
// adjustment: SequentialActuator adds one image to the map and immediately runs a full BA over everything, every time
// (src/actuator/SequentialActuator.h:138-229, the call at :227-228 constructs a fresh BundleAdjuster, which rebuilds the whole
// ceres::Problem from the map, BundleAdjuster.h:100-121).  This class is meant to LIVE across those calls (a member of the actuator
// instead of a temporary): it keeps one engine handle, and when the map has only GROWN since the previous call it sends just the
// new poses, points and observations (ba_append) — the parameters of everything older are already on the device, where the previous
// solve left them.  Anything else (a point or image removed, an observation dropped or edited, the constant pose changed) makes
// it fall back to a full ba_set_problem, so the result is always what BundleAdjuster would produce on the same world:
//   * poses and intrinsics are small and go through Image/Camera conversions on write-back, so they are always refreshed from the
//     world (ba_update_parameters) — this also reproduces the reference's default-inserted all-zero first pose (F8b) on every call;
//   * points are compared with what the last call wrote back and refreshed only where the world edited them.
// Same failure path as BundleAdjuster: an unusable solve leaves the world untouched (:128-131,180-184).
#ifndef B200BA_INCREMENTAL_BUNDLE_ADJUSTER_H
#define B200BA_INCREMENTAL_BUNDLE_ADJUSTER_H

#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "../../../include/ba_engine.h"
#include "../world/WorldStructure.h"
#include "WorldGather.h"

class IncrementalBundleAdjuster {
    ba_options options_{};
    ba_engine* engine_ = nullptr;
    WorldGather held_;                  // what the handle holds: identities, structure, and the state written back last time
    ba_summary summary_{};
    std::vector<ba_iter_record> trajectory_;
    int appended_calls_ = 0, rebuilt_calls_ = 0;

    static ba_problem asProblem(const WorldGather& g) {
        ba_problem p{};
        p.num_cameras = int32_t(g.images_.size()); p.num_groups = int32_t(g.cameras_.size());
        p.num_points = int32_t(g.points_.size()); p.num_observations = int32_t(g.obs_cam_.size());
        p.pose = g.pose_.data(); p.intr = g.intr_.data(); p.cam_group = g.cam_group_.data(); p.cam_fixed = g.cam_fixed_.data();
        p.pt = g.pt_.data(); p.obs_cam = g.obs_cam_.data(); p.obs_pt = g.obs_pt_.data(); p.obs_uv = g.obs_uv_.data();
        return p;
    }

    // true when `now` is `held_` plus new images / cameras / points at the end and new observations at the end of each point's list
    bool onlyGrown(const WorldGather& now, std::vector<int32_t>& new_obs) const {
        const size_t Nc = held_.images_.size(), Ng = held_.cameras_.size(), Np = held_.points_.size();
        if (now.images_.size() < Nc || now.cameras_.size() < Ng || now.points_.size() < Np) return false;
        for (size_t c = 0; c < Nc; ++c)
            if (now.images_[c] != held_.images_[c] || now.cam_group_[c] != held_.cam_group_[c] || now.cam_fixed_[c] != held_.cam_fixed_[c]) return false;
        for (size_t g = 0; g < Ng; ++g) if (now.cameras_[g] != held_.cameras_[g]) return false;
        for (size_t p = 0; p < Np; ++p) if (now.points_[p] != held_.points_[p]) return false;
        // both observation lists are sorted by point: walk them together
        size_t i = 0, j = 0;
        const size_t no = held_.obs_cam_.size(), nn = now.obs_cam_.size();
        while (j < nn) {
            const int32_t p = now.obs_pt_[j];
            while (j < nn && now.obs_pt_[j] == p) {
                if (i < no && held_.obs_pt_[i] == p) {
                    if (held_.obs_cam_[i] != now.obs_cam_[j] || held_.obs_uv_[2 * i] != now.obs_uv_[2 * j] || held_.obs_uv_[2 * i + 1] != now.obs_uv_[2 * j + 1]) return false;
                    ++i;
                } else {
                    new_obs.push_back(int32_t(j));
                }
                ++j;
            }
            if (i < no && held_.obs_pt_[i] == p) return false;   // the point lost observations
            if (i < no && held_.obs_pt_[i] < p) return false;    // a point in between lost all of them
        }
        return i == no;
    }

    int sendGrowth(const WorldGather& now, const std::vector<int32_t>& new_obs) {
        const size_t Nc = held_.images_.size(), Ng = held_.cameras_.size(), Np = held_.points_.size();
        std::vector<int32_t> oc, op;
        std::vector<double> uv;
        for (int32_t j : new_obs) { oc.push_back(now.obs_cam_[j]); op.push_back(now.obs_pt_[j]); uv.push_back(now.obs_uv_[2 * size_t(j)]); uv.push_back(now.obs_uv_[2 * size_t(j) + 1]); }
        ba_delta d{};
        d.num_new_cameras = int32_t(now.images_.size() - Nc); d.pose = now.pose_.data() + 6 * Nc; d.cam_group = now.cam_group_.data() + Nc; d.cam_fixed = now.cam_fixed_.data() + Nc;
        d.num_new_groups = int32_t(now.cameras_.size() - Ng); d.intr = now.intr_.data() + 4 * Ng;
        d.num_new_points = int32_t(now.points_.size() - Np); d.pt = now.pt_.data() + 3 * Np;
        d.num_new_observations = int32_t(oc.size()); d.obs_cam = oc.data(); d.obs_pt = op.data(); d.obs_uv = uv.data();
        int rc = ba_append(engine_, &d);
        // refresh what the world may have edited since the last write-back: all poses and intrinsics (small), changed points only
        if (rc == BA_OK && Nc) rc = ba_update_parameters(engine_, 0, 0, int32_t(Nc), now.pose_.data());
        if (rc == BA_OK && Ng) rc = ba_update_parameters(engine_, 1, 0, int32_t(Ng), now.intr_.data());
        for (size_t p = 0; rc == BA_OK && p < Np;) {
            if (std::memcmp(&now.pt_[3 * p], &held_.pt_[3 * p], 3 * sizeof(double)) == 0) { ++p; continue; }
            size_t q = p + 1;
            while (q < Np && std::memcmp(&now.pt_[3 * q], &held_.pt_[3 * q], 3 * sizeof(double)) != 0) ++q;
            rc = ba_update_parameters(engine_, 2, int32_t(p), int32_t(q - p), &now.pt_[3 * p]);
            p = q;
        }
        return rc;
    }

public:
    [[nodiscard]] IncrementalBundleAdjuster() { ba_options_default(&options_); }
    IncrementalBundleAdjuster(const IncrementalBundleAdjuster&) = delete;
    IncrementalBundleAdjuster& operator=(const IncrementalBundleAdjuster&) = delete;
    ~IncrementalBundleAdjuster() { if (engine_) ba_destroy(engine_); }

    ba_options& options() { return options_; }             // read when the handle is created (first call)
    const ba_summary& summary() const { return summary_; }
    const std::vector<ba_iter_record>& trajectory() const { return trajectory_; }
    int appendedCalls() const { return appended_calls_; }  // calls served by ba_append
    int rebuiltCalls() const { return rebuilt_calls_; }    // calls that needed a full ba_set_problem

    void operator()(WorldStructure::Ptr& world) {
        WorldGather now;
        now.load(world->world_points_, world->local_frames_);
        int rc = BA_OK;
        if (!engine_) {
            rc = ba_create(&options_, &engine_);
            if (rc != BA_OK) { std::cout << "Bundle Adjustment failed. (" << ba_last_error(nullptr) << ")" << std::endl; engine_ = nullptr; return; }
        }
        std::vector<int32_t> new_obs;
        const bool grow = !held_.images_.empty() && onlyGrown(now, new_obs);
        if (grow) { rc = sendGrowth(now, new_obs); ++appended_calls_; }
        else { const ba_problem p = asProblem(now); rc = ba_set_problem(engine_, &p); ++rebuilt_calls_; }
        if (rc == BA_OK) rc = ba_solve(engine_, &summary_);
        if (rc == BA_OK) {
            int32_t n = 0;
            ba_get_trajectory(engine_, nullptr, 0, &n);
            trajectory_.resize(size_t(n));
            if (n) ba_get_trajectory(engine_, trajectory_.data(), n, &n);
        }
        if (rc != BA_OK || !summary_.usable) {
            std::cout << "Bundle Adjustment failed." << (rc != BA_OK ? std::string(" (") + ba_last_error(engine_) + ")" : std::string()) << std::endl;
            held_.clear();   // the handle no longer matches the world: next call starts over
            return;
        }
        rc = ba_get_state(engine_, now.pose_.data(), now.intr_.data(), now.pt_.data());
        if (rc != BA_OK) { std::cout << "Bundle Adjustment failed." << std::endl; held_.clear(); return; }
        std::cout << "Bundle Adjustment statistics (approximated RMSE):\n"
                  << "    #views: " << now.images_.size() << "\n"
                  << "    #residuals: " << summary_.num_residuals << "\n"
                  << "    Initial RMSE: " << std::sqrt(summary_.initial_cost / summary_.num_residuals) << "\n"
                  << "    Final RMSE: " << std::sqrt(summary_.final_cost / summary_.num_residuals) << "\n"
                  << "    Time (s): " << summary_.solve_seconds << "\n";
        now.store();
        held_ = std::move(now);
    }
};

#endif  // B200BA_INCREMENTAL_BUNDLE_ADJUSTER_H
