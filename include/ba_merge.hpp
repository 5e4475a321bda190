// ba_merge.hpp — host-side helper shared by the engine and the CPU oracle: validate a ba_delta against the problem a handle holds
// and merge it into canonical order (observations sorted by point; within a point the existing observations first, then the new
// ones in the order given).  Header-only C++, no device code.  (ba_append, include/ba_engine.h)
#pragma once
#include <algorithm>
#include <cstdint>
#include <numeric>
#include <string>
#include <vector>

#include "ba_engine.h"

namespace ba_merge {

struct Merged {
    std::vector<int32_t> cam_group, obs_cam, obs_pt;
    std::vector<uint8_t> cam_fixed;
    std::vector<double> obs_uv;
    int Nc = 0, Ng = 0, Np = 0, No = 0;
};

// Returns an empty string on success, else what is wrong with the delta.
inline std::string merge(int Nc, int Ng, int Np, int No, const int32_t* cam_group, const uint8_t* cam_fixed, const int32_t* obs_cam,
                         const int32_t* obs_pt, const double* obs_uv, const ba_delta& d, Merged& M) {
    if (d.num_new_cameras < 0 || d.num_new_groups < 0 || d.num_new_points < 0 || d.num_new_observations < 0) return "negative size";
    if ((d.num_new_cameras && (!d.pose || !d.cam_group || !d.cam_fixed)) || (d.num_new_groups && !d.intr) || (d.num_new_points && !d.pt) ||
        (d.num_new_observations && (!d.obs_cam || !d.obs_pt || !d.obs_uv)))
        return "null array";
    const int64_t no64 = int64_t(No) + d.num_new_observations;
    if (no64 > INT32_MAX || int64_t(Nc) + d.num_new_cameras > INT32_MAX || int64_t(Np) + d.num_new_points > INT32_MAX) return "problem too large";
    M.Nc = Nc + d.num_new_cameras; M.Ng = Ng + d.num_new_groups; M.Np = Np + d.num_new_points; M.No = int(no64);
    for (int c = 0; c < d.num_new_cameras; ++c)
        if (d.cam_group[c] < 0 || d.cam_group[c] >= M.Ng) return "cam_group out of range";
    for (int o = 0; o < d.num_new_observations; ++o)
        if (d.obs_cam[o] < 0 || d.obs_cam[o] >= M.Nc || d.obs_pt[o] < 0 || d.obs_pt[o] >= M.Np) return "observation index out of range";
    M.cam_group.assign(cam_group, cam_group + Nc); M.cam_group.insert(M.cam_group.end(), d.cam_group, d.cam_group + d.num_new_cameras);
    M.cam_fixed.assign(cam_fixed, cam_fixed + Nc); M.cam_fixed.insert(M.cam_fixed.end(), d.cam_fixed, d.cam_fixed + d.num_new_cameras);
    // new observations in point order (stable), then a two-way merge with the existing (already sorted) list
    std::vector<int32_t> idx(d.num_new_observations);
    std::iota(idx.begin(), idx.end(), 0);
    std::stable_sort(idx.begin(), idx.end(), [&](int32_t a, int32_t b) { return d.obs_pt[a] < d.obs_pt[b]; });
    M.obs_cam.resize(M.No); M.obs_pt.resize(M.No); M.obs_uv.resize(size_t(2) * M.No);
    int i = 0, j = 0, k = 0;
    while (i < No || j < d.num_new_observations) {
        const bool take_old = j >= d.num_new_observations || (i < No && obs_pt[i] <= d.obs_pt[idx[j]]);
        if (take_old) { M.obs_cam[k] = obs_cam[i]; M.obs_pt[k] = obs_pt[i]; M.obs_uv[2 * size_t(k)] = obs_uv[2 * size_t(i)]; M.obs_uv[2 * size_t(k) + 1] = obs_uv[2 * size_t(i) + 1]; ++i; }
        else { const int q = idx[j]; M.obs_cam[k] = d.obs_cam[q]; M.obs_pt[k] = d.obs_pt[q]; M.obs_uv[2 * size_t(k)] = d.obs_uv[2 * size_t(q)]; M.obs_uv[2 * size_t(k) + 1] = d.obs_uv[2 * size_t(q) + 1]; ++j; }
        ++k;
    }
    return {};
}

}  // namespace ba_merge
