// ba_bal.cpp — reader / writer for the "Bundle Adjustment in the Large" text format, the data format of the model the reference
// keeps in src/adjuster/SnavelyReprojectionError.h:16-55 (camera = angle-axis, t, f, k1, k2; -z forward).  Host only.
//   <num_cameras> <num_points> <num_observations>
//   <camera> <point> <x> <y>                    x num_observations
//   9 lines per camera: r0 r1 r2 t0 t1 t2 f k1 k2
//   3 lines per point
// Output is a ba_synth_problem in the canonical indexing of include/ba_engine.h: observations stably sorted by point id, one
// intrinsics group per camera holding (f, k1, k2, 0), nothing fixed.  Solve it with ba_options.model = BA_MODEL_SNAVELY.
#include "../../include/ba_synth.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <vector>

extern "C" int ba_bal_read(const char* path, ba_synth_problem** out) {
    if (!path || !out) return -1;
    FILE* f = std::fopen(path, "r");
    if (!f) return -2;
    int Nc = 0, Np = 0, No = 0;
    if (std::fscanf(f, "%d %d %d", &Nc, &Np, &No) != 3 || Nc < 0 || Np < 0 || No < 0) { std::fclose(f); return -3; }
    std::vector<int> cam(No), pt(No);
    std::vector<double> uv(size_t(2) * No);
    for (int o = 0; o < No; ++o)
        if (std::fscanf(f, "%d %d %lf %lf", &cam[o], &pt[o], &uv[2 * o], &uv[2 * o + 1]) != 4 || cam[o] < 0 || cam[o] >= Nc || pt[o] < 0 || pt[o] >= Np) { std::fclose(f); return -3; }
    std::vector<double> cams(size_t(9) * Nc), pts(size_t(3) * Np);
    for (double& v : cams) if (std::fscanf(f, "%lf", &v) != 1) { std::fclose(f); return -3; }
    for (double& v : pts) if (std::fscanf(f, "%lf", &v) != 1) { std::fclose(f); return -3; }
    std::fclose(f);
    std::vector<int> order(No);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return pt[a] < pt[b]; });

    ba_synth_problem* P = static_cast<ba_synth_problem*>(std::calloc(1, sizeof(ba_synth_problem)));
    if (!P) return -4;
    P->num_cameras = Nc; P->num_groups = Nc; P->num_points = Np; P->num_observations = No;
    auto alloc = [](size_t n, size_t sz) { return std::calloc(std::max<size_t>(n, 1), sz); };
    P->pose = static_cast<double*>(alloc(size_t(6) * Nc, 8)); P->intr = static_cast<double*>(alloc(size_t(4) * Nc, 8));
    P->cam_group = static_cast<int32_t*>(alloc(Nc, 4)); P->cam_fixed = static_cast<uint8_t*>(alloc(Nc, 1));
    P->pt = static_cast<double*>(alloc(size_t(3) * Np, 8));
    P->obs_cam = static_cast<int32_t*>(alloc(No, 4)); P->obs_pt = static_cast<int32_t*>(alloc(No, 4)); P->obs_uv = static_cast<double*>(alloc(size_t(2) * No, 8));
    P->pose_true = static_cast<double*>(alloc(size_t(6) * Nc, 8)); P->pt_true = static_cast<double*>(alloc(size_t(3) * Np, 8));
    for (int c = 0; c < Nc; ++c) {
        for (int k = 0; k < 6; ++k) P->pose[6 * c + k] = P->pose_true[6 * c + k] = cams[9 * c + k];
        for (int k = 0; k < 3; ++k) P->intr[4 * c + k] = cams[9 * c + 6 + k];
        P->cam_group[c] = c;
    }
    for (size_t i = 0; i < pts.size(); ++i) P->pt[i] = P->pt_true[i] = pts[i];
    for (int i = 0; i < No; ++i) { const int o = order[i]; P->obs_cam[i] = cam[o]; P->obs_pt[i] = pt[o]; P->obs_uv[2 * i] = uv[2 * o]; P->obs_uv[2 * i + 1] = uv[2 * o + 1]; }
    *out = P;
    return 0;
}

extern "C" int ba_bal_write(const char* path, int32_t Nc, int32_t Np, int32_t No, const double* pose, const double* intr, const int32_t* cam_group,
                            const double* pt, const int32_t* obs_cam, const int32_t* obs_pt, const double* obs_uv) {
    if (!path) return -1;
    FILE* f = std::fopen(path, "w");
    if (!f) return -2;
    std::fprintf(f, "%d %d %d\n", Nc, Np, No);
    for (int o = 0; o < No; ++o) std::fprintf(f, "%d %d %.17g %.17g\n", obs_cam[o], obs_pt[o], obs_uv[2 * o], obs_uv[2 * o + 1]);
    for (int c = 0; c < Nc; ++c) {
        for (int k = 0; k < 6; ++k) std::fprintf(f, "%.17g\n", pose[6 * c + k]);
        const int g = cam_group ? cam_group[c] : c;
        for (int k = 0; k < 3; ++k) std::fprintf(f, "%.17g\n", intr[4 * g + k]);
    }
    for (int i = 0; i < 3 * Np; ++i) std::fprintf(f, "%.17g\n", pt[i]);
    std::fclose(f);
    return 0;
}

// ASCII PLY of the map points plus the camera centres C = -R^T t (green), as the reference's exportToPly writes them
// (src/helpers/SfMPlyHelper.hpp:62-116; WorldStructure::writeToPLYFile, src/world/WorldStructure.h:83-98).  No viewer: graphics
// are unavailable on this box, a file is all there is.
extern "C" int ba_write_ply(const char* path, int32_t Np, const double* pt, int32_t Nc, const double* pose) {
    if (!path) return -1;
    FILE* f = std::fopen(path, "w");
    if (!f) return -2;
    std::fprintf(f, "ply\nformat ascii 1.0\nelement vertex %d\nproperty double x\nproperty double y\nproperty double z\n"
                    "property uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n", Np + Nc);
    for (int p = 0; p < Np; ++p) std::fprintf(f, "%.17g %.17g %.17g 255 255 255\n", pt[3 * p], pt[3 * p + 1], pt[3 * p + 2]);
    for (int c = 0; c < Nc; ++c) {
        const double* a = pose + 6 * c;
        const double th2 = a[0] * a[0] + a[1] * a[1] + a[2] * a[2];
        // C = -R(aa)^T t = -R(-aa) t  (Rodrigues with the negated axis)
        double C[3];
        const double w[3] = {-a[0], -a[1], -a[2]}, t[3] = {a[3], a[4], a[5]};
        if (th2 > 0.0) {
            const double th = std::sqrt(th2), cs = std::cos(th), sn = std::sin(th);
            const double k[3] = {w[0] / th, w[1] / th, w[2] / th};
            const double kxt[3] = {k[1] * t[2] - k[2] * t[1], k[2] * t[0] - k[0] * t[2], k[0] * t[1] - k[1] * t[0]};
            const double kt = k[0] * t[0] + k[1] * t[1] + k[2] * t[2];
            for (int i = 0; i < 3; ++i) C[i] = -(t[i] * cs + kxt[i] * sn + k[i] * kt * (1.0 - cs));
        } else for (int i = 0; i < 3; ++i) C[i] = -t[i];
        std::fprintf(f, "%.17g %.17g %.17g 0 255 0\n", C[0], C[1], C[2]);
    }
    std::fclose(f);
    return 0;
}
