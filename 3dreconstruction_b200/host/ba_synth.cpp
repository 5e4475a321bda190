// ba_synth.cpp — deterministic synthetic BAL-shape scenes (see include/ba_synth.h).
// Host only. std::mt19937_64 is bit-exact across standard libraries; uniforms and normals are
// derived by hand (std::*_distribution is implementation-defined).
#include "../../include/ba_synth.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

namespace {

struct Rng {
    std::mt19937_64 g;
    explicit Rng(uint64_t seed) : g(seed) {}
    double uniform() { return double(g() >> 11) * (1.0 / 9007199254740992.0); }  // [0,1)
    double normal() {                                                              // Box-Muller, one value per call
        double u1 = uniform(), u2 = uniform();
        if (u1 < 1e-300) u1 = 1e-300;
        return std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586476925 * u2);
    }
    int32_t below(int32_t n) { return int32_t(g() % uint64_t(n)); }
};

struct Vec3 { double x, y, z; };
inline Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator*(double s, Vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline double dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vec3 cross(Vec3 a, Vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline Vec3 normalized(Vec3 a) { double n = std::sqrt(dot(a, a)); return {a.x / n, a.y / n, a.z / n}; }

struct Pose { double R[9]; double t[3]; };  // world -> camera, row-major R

// Rotation matrix -> angle-axis through the unit quaternion (stable for angles up to pi).
void rotation_to_angle_axis(const double* R, double* aa) {
    double q[4];  // w x y z
    double tr = R[0] + R[4] + R[8];
    if (tr >= 0.0) {
        double t = std::sqrt(tr + 1.0);
        q[0] = 0.5 * t; t = 0.5 / t;
        q[1] = (R[7] - R[5]) * t; q[2] = (R[2] - R[6]) * t; q[3] = (R[3] - R[1]) * t;
    } else {
        int i = 0;
        if (R[4] > R[0]) i = 1;
        if (R[8] > R[i * 4]) i = 2;
        int j = (i + 1) % 3, k = (j + 1) % 3;
        double t = std::sqrt(R[i * 4] - R[j * 4] - R[k * 4] + 1.0);
        q[i + 1] = 0.5 * t; t = 0.5 / t;
        q[0] = (R[k * 3 + j] - R[j * 3 + k]) * t;
        q[j + 1] = (R[j * 3 + i] + R[i * 3 + j]) * t;
        q[k + 1] = (R[k * 3 + i] + R[i * 3 + k]) * t;
    }
    if (q[0] < 0) { for (double& v : q) v = -v; }
    double s = std::sqrt(q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    if (s < 1e-300) { aa[0] = aa[1] = aa[2] = 0.0; return; }
    double k = 2.0 * std::atan2(s, q[0]) / s;
    aa[0] = q[1] * k; aa[1] = q[2] * k; aa[2] = q[3] * k;
}

void angle_axis_to_rotation(const double* aa, double* R) {
    double th = std::sqrt(aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2]);
    if (th < 1e-300) { double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; std::memcpy(R, I, sizeof I); return; }
    double c = std::cos(th), s = std::sin(th), x = aa[0] / th, y = aa[1] / th, z = aa[2] / th, C = 1 - c;
    R[0] = c + x * x * C;     R[1] = x * y * C - z * s; R[2] = x * z * C + y * s;
    R[3] = y * x * C + z * s; R[4] = c + y * y * C;     R[5] = y * z * C - x * s;
    R[6] = z * x * C - y * s; R[7] = z * y * C + x * s; R[8] = c + z * z * C;
}

inline void project(const double* R, const double* t, const double* X, const ba_synth_spec& s, double* uv, double* depth) {
    double P0 = R[0] * X[0] + R[1] * X[1] + R[2] * X[2] + t[0];
    double P1 = R[3] * X[0] + R[4] * X[1] + R[5] * X[2] + t[1];
    double P2 = R[6] * X[0] + R[7] * X[1] + R[8] * X[2] + t[2];
    uv[0] = s.fx * P0 / P2 + s.cx; uv[1] = s.fy * P1 / P2 + s.cy; *depth = P2;
}

template <class T> T* dup(const std::vector<T>& v) {
    T* p = static_cast<T*>(std::malloc(std::max<size_t>(1, v.size()) * sizeof(T)));
    if (p && !v.empty()) std::memcpy(p, v.data(), v.size() * sizeof(T));
    return p;
}

}  // namespace

extern "C" int ba_synth_spec_default(ba_synth_spec* s) {
    if (!s) return -1;
    std::memset(s, 0, sizeof *s);
    s->num_cameras = 6; s->num_points = 40; s->num_observations = 140; s->seed = 0xBA5E0000ull;
    s->max_track = 40;
    s->fx = 2905.88; s->fy = 2905.88; s->cx = 1416; s->cy = 1064;
    s->noise_px = 0.5; s->outlier_fraction = 0.02; s->outlier_px = 30.0;
    s->sigma_aa = 0.01; s->sigma_t_rel = 0.01; s->sigma_pt_rel = 0.02;
    s->round_obs_to_float = 0; s->first_pose_zero = 0; s->fixed_camera = 0; s->arc = 0;
    return 0;
}

extern "C" int ba_synth_named(const char* name, ba_synth_spec* s) {
    if (!name || !s) return -1;
    ba_synth_spec_default(s);
    struct Row { const char* n; int32_t nc, np, no, id; };
    static const Row rows[] = {{"tiny", 6, 40, 140, 0},           {"ladybug", 49, 7776, 31843, 2},
                               {"trafalgar", 257, 65132, 225911, 3}, {"venice", 1778, 993923, 5001946, 4},
                               {"final", 13682, 4456117, 28987644, 5}, {"sequential30", 30, 7250, 25000, 1}};
    for (const Row& r : rows)
        if (std::strcmp(name, r.n) == 0) {
            s->num_cameras = r.nc; s->num_points = r.np; s->num_observations = r.no; s->seed = 0xBA5E0000ull + uint64_t(r.id);
            if (r.id == 1) {  // what SequentialActuator hands to BA
                s->round_obs_to_float = 1; s->first_pose_zero = 1; s->fixed_camera = 1; s->arc = 1; s->max_track = 6;
            }
            return 0;
        }
    return -1;
}

extern "C" int ba_synth_generate(const ba_synth_spec* sp, ba_synth_problem** out) {
    if (!sp || !out) return -1;
    const ba_synth_spec& s = *sp;
    const int32_t Nc = s.num_cameras, Np = s.num_points, No = s.num_observations;
    const int32_t kmax = std::min(Nc, s.max_track);
    if (Nc < 2 || Np < 1 || kmax < 2 || int64_t(No) < 2 * int64_t(Np) || int64_t(No) > int64_t(Np) * kmax) return -2;
    Rng rng(s.seed);

    // ---- scene: points in a box, cameras on a ring (or a 120 degree arc) looking at the centre ----
    const double sigma = std::cbrt(double(Np)) / 20.0;
    const double extent = 10.0 * sigma;
    const double radius = 4.5 * extent;
    std::vector<Pose> cam(Nc);
    for (int32_t i = 0; i < Nc; ++i) {
        double phi = s.arc ? ((Nc > 1 ? double(i) / (Nc - 1) : 0.5) - 0.5) * (2.0943951023931953)
                           : 6.283185307179586476925 * double(i) / Nc;
        double rr = radius * (1.0 + 0.05 * (2.0 * rng.uniform() - 1.0));
        double h = 0.3 * extent * (2.0 * rng.uniform() - 1.0);
        Vec3 C{rr * std::sin(phi), h, -rr * std::cos(phi)};
        Vec3 f = normalized(Vec3{-C.x, -C.y, -C.z});          // +z forward (ReprojectCost convention, BundleAdjuster.h:52-53)
        Vec3 yw{0, 1, 0};
        Vec3 d = normalized(yw - dot(yw, f) * f);             // image "down"
        Vec3 r = cross(d, f);
        double R[9] = {r.x, r.y, r.z, d.x, d.y, d.z, f.x, f.y, f.z};
        std::memcpy(cam[i].R, R, sizeof R);
        for (int a = 0; a < 3; ++a) cam[i].t[a] = -(R[a * 3] * C.x + R[a * 3 + 1] * C.y + R[a * 3 + 2] * C.z);
    }

    // ---- track lengths: 2 + geometric, truncated to [2,kmax], then nudged so that sum == No exactly ----
    std::vector<int32_t> klen(Np);
    const double mean = double(No) / double(Np);
    const double q = 1.0 / std::max(1.0 + 1e-9, mean - 1.0);  // success probability, E[G] = mean-2
    int64_t total = 0;
    for (int32_t p = 0; p < Np; ++p) {
        int32_t k = 2;
        if (q < 1.0) {
            double u = rng.uniform();
            k = 2 + int32_t(std::floor(std::log(1.0 - u) / std::log(1.0 - q)));
        }
        klen[p] = std::min(std::max(k, 2), kmax);
        total += klen[p];
    }
    while (total != No) {
        int32_t p = rng.below(Np);
        if (total < No && klen[p] < kmax) { ++klen[p]; ++total; }
        else if (total > No && klen[p] > 2) { --klen[p]; --total; }
    }

    // ---- points and visibility (generation order; renumbered by birth camera below) ----
    std::vector<double> X(size_t(Np) * 3);
    std::vector<int64_t> off(size_t(Np) + 1, 0);
    for (int32_t p = 0; p < Np; ++p) off[p + 1] = off[p] + klen[p];
    std::vector<int32_t> ocam(No);
    std::vector<int32_t> birth(Np);
    const int32_t stride = std::max(1, Nc / 128);
    std::vector<int32_t> pick;
    for (int32_t p = 0; p < Np; ++p) {
        X[p * 3 + 0] = extent * (2.0 * rng.uniform() - 1.0);
        X[p * 3 + 1] = 0.3 * extent * (2.0 * rng.uniform() - 1.0);
        X[p * 3 + 2] = extent * (2.0 * rng.uniform() - 1.0);
        const int32_t k = klen[p];
        int32_t nwin = std::min(k, int32_t(std::ceil(0.8 * k)));
        int32_t st = stride;
        while (st > 1 && int64_t(nwin) * st > Nc) --st;
        const int32_t centre = rng.below(Nc);
        pick.clear();
        int32_t start = centre - (nwin / 2) * st;
        if (s.arc) start = std::min(std::max(start, 0), Nc - 1 - (nwin - 1) * st);
        for (int32_t j = 0; j < nwin; ++j) pick.push_back(((start + j * st) % Nc + Nc) % Nc);   // strided window: sequence locality
        while (int32_t(pick.size()) < k) {                                                      // the rest anywhere: loop closures
            int32_t c = rng.below(Nc);
            if (std::find(pick.begin(), pick.end(), c) != pick.end()) continue;
            pick.push_back(c);
        }
        std::sort(pick.begin(), pick.end());
        birth[p] = pick[0];
        std::copy(pick.begin(), pick.end(), ocam.begin() + off[p]);
    }

    // ---- renumber points by birth camera (stable counting sort) ----
    std::vector<int32_t> order(Np);
    {
        std::vector<int32_t> cnt(size_t(Nc) + 1, 0);
        for (int32_t p = 0; p < Np; ++p) ++cnt[birth[p] + 1];
        for (int32_t c = 0; c < Nc; ++c) cnt[c + 1] += cnt[c];
        for (int32_t p = 0; p < Np; ++p) order[cnt[birth[p]]++] = p;
    }

    std::vector<double> pt_true(size_t(Np) * 3), uv(size_t(No) * 2);
    std::vector<int32_t> obs_cam(No), obs_pt(No);
    {
        int64_t o = 0;
        for (int32_t n = 0; n < Np; ++n) {
            const int32_t p = order[n];
            for (int a = 0; a < 3; ++a) pt_true[size_t(n) * 3 + a] = X[size_t(p) * 3 + a];
            for (int64_t j = off[p]; j < off[p + 1]; ++j, ++o) { obs_cam[o] = ocam[j]; obs_pt[o] = n; }
        }
    }

    // ---- optional change of world frame so that camera 0 is exactly the identity (reference flow, F8b) ----
    if (s.first_pose_zero) {
        const Pose c0 = cam[0];
        for (int32_t n = 0; n < Np; ++n) {
            double* Xn = &pt_true[size_t(n) * 3];
            double Y[3];
            for (int a = 0; a < 3; ++a) Y[a] = c0.R[a * 3] * Xn[0] + c0.R[a * 3 + 1] * Xn[1] + c0.R[a * 3 + 2] * Xn[2] + c0.t[a];
            Xn[0] = Y[0]; Xn[1] = Y[1]; Xn[2] = Y[2];
        }
        for (int32_t i = 0; i < Nc; ++i) {
            double Rn[9];
            for (int a = 0; a < 3; ++a)
                for (int b = 0; b < 3; ++b)   // R_i * R_0^T
                    Rn[a * 3 + b] = cam[i].R[a * 3] * c0.R[b * 3] + cam[i].R[a * 3 + 1] * c0.R[b * 3 + 1] + cam[i].R[a * 3 + 2] * c0.R[b * 3 + 2];
            double tn[3];
            for (int a = 0; a < 3; ++a) tn[a] = cam[i].t[a] - (Rn[a * 3] * c0.t[0] + Rn[a * 3 + 1] * c0.t[1] + Rn[a * 3 + 2] * c0.t[2]);
            std::memcpy(cam[i].R, Rn, sizeof Rn); std::memcpy(cam[i].t, tn, sizeof tn);
        }
        double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        std::memcpy(cam[0].R, I, sizeof I); cam[0].t[0] = cam[0].t[1] = cam[0].t[2] = 0.0;
    }

    // ---- ground-truth poses as (angle-axis, t); re-derive R from the rounded angle-axis so truth is self-consistent ----
    std::vector<double> pose_true(size_t(Nc) * 6);
    for (int32_t i = 0; i < Nc; ++i) {
        rotation_to_angle_axis(cam[i].R, &pose_true[size_t(i) * 6]);
        for (int a = 0; a < 3; ++a) pose_true[size_t(i) * 6 + 3 + a] = cam[i].t[a];
        angle_axis_to_rotation(&pose_true[size_t(i) * 6], cam[i].R);
    }

    // ---- observations: exact projection + Gaussian noise (+ sparse gross outliers) ----
    for (int64_t o = 0; o < No; ++o) {
        double depth;
        project(cam[obs_cam[o]].R, cam[obs_cam[o]].t, &pt_true[size_t(obs_pt[o]) * 3], s, &uv[o * 2], &depth);
        uv[o * 2] += s.noise_px * rng.normal();
        uv[o * 2 + 1] += s.noise_px * rng.normal();
        if (s.outlier_fraction > 0 && rng.uniform() < s.outlier_fraction) {
            uv[o * 2] += s.outlier_px * (2.0 * rng.uniform() - 1.0);
            uv[o * 2 + 1] += s.outlier_px * (2.0 * rng.uniform() - 1.0);
        }
        if (s.round_obs_to_float) { uv[o * 2] = double(float(uv[o * 2])); uv[o * 2 + 1] = double(float(uv[o * 2 + 1])); }
    }

    // ---- initial guess = truth + perturbation; the fixed pose (and an exactly-zero pose 0) stay at truth ----
    std::vector<double> pose(pose_true), pt(pt_true);
    std::vector<uint8_t> fixed(Nc, 0);
    if (s.fixed_camera >= 0 && s.fixed_camera < Nc) fixed[s.fixed_camera] = 1;
    for (int32_t i = 0; i < Nc; ++i) {
        double d[6];
        for (int a = 0; a < 3; ++a) d[a] = s.sigma_aa * rng.normal();
        for (int a = 3; a < 6; ++a) d[a] = s.sigma_t_rel * extent * rng.normal();
        if (fixed[i] || (s.first_pose_zero && i == 0)) continue;
        for (int a = 0; a < 6; ++a) pose[size_t(i) * 6 + a] += d[a];
    }
    for (size_t j = 0; j < pt.size(); ++j) pt[j] += s.sigma_pt_rel * extent * rng.normal();

    ba_synth_problem* P = static_cast<ba_synth_problem*>(std::calloc(1, sizeof(ba_synth_problem)));
    if (!P) return -3;
    P->num_cameras = Nc; P->num_groups = 1; P->num_points = Np; P->num_observations = No; P->extent = extent;
    std::vector<double> intr{s.fx, s.fy, s.cx, s.cy};
    std::vector<int32_t> group(Nc, 0);
    P->pose = dup(pose); P->intr = dup(intr); P->cam_group = dup(group); P->cam_fixed = dup(fixed); P->pt = dup(pt);
    P->obs_cam = dup(obs_cam); P->obs_pt = dup(obs_pt); P->obs_uv = dup(uv); P->pose_true = dup(pose_true); P->pt_true = dup(pt_true);
    *out = P;
    return 0;
}

extern "C" void ba_synth_free(ba_synth_problem* p) {
    if (!p) return;
    std::free(p->pose); std::free(p->intr); std::free(p->cam_group); std::free(p->cam_fixed); std::free(p->pt);
    std::free(p->obs_cam); std::free(p->obs_pt); std::free(p->obs_uv); std::free(p->pose_true); std::free(p->pt_true);
    std::free(p);
}
