// oracle/ba_oracle.cpp — CPU oracle: a dependency-free restatement of the reference's bundle-adjustment path
// (src/adjuster/BundleAdjuster.h) behind the same C-ABI as the CUDA engine (include/ba_engine.h).
//
// TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load this library; the product path (3dreconstruction_b200/) never links, imports or executes it.
//
// PARITY: PINNED FOR THE RESIDUAL MODELS, UNPINNED FOR THE SOLVER.  The reference's own cost functors (ReprojectCost,
// SnavelyReprojectionError) are compiled here from /root/reference by oracle/ref_build.sh (oracle/_ref/libref_cost.so) and this
// oracle's residuals and Jacobians equal theirs to 1e-13 (tests/test_reference_functors.py, vectors in tests/golden/).  Beyond that the
// reference holds no golden vectors, known-answer tests or fixtures for this path
//  and cannot be built here (F5: Ceres 2.2, Eigen, OpenCV, Sophus, PCL absent).  The solver the
// reference delegates to — Ceres Solver 2.2 (CMakeLists.txt:32, un-vendored, patch version unpinned),
// TrustRegionMinimizer + LevenbergMarquardtStrategy + SPARSE_SCHUR/EIGEN_SPARSE as configured at
// BundleAdjuster.h:167-174 — is restated below from its published algorithm (SURVEY §8 a-LM) and anchored on the
// reference's call sites: problem construction :100-121, Solve :126, usability test :128, gather/scatter
// :82-98 / :143-156.  Independent cross-checks (OpenCV, SciPy, finite differences) are in tests/.
//
// What follows the reference line by line:
//   parameter blocks / fixed block / shared loss / residual order   BundleAdjuster.h:102-121  -> set_problem(), evaluate()
//   residual + autodiff Jacobian                                     BundleAdjuster.h:38-68,113 -> residuals.h + jet.h
//   solver options                                                   BundleAdjuster.h:167-174 -> ba_options_default()
#include "../include/ba_engine.h"
#include "../include/ba_merge.hpp"

#include <algorithm>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "residuals.h"

namespace {
using std::vector;
thread_local std::string g_create_error;

inline bool finite_all(const double* p, int n) { for (int i = 0; i < n; ++i) if (!std::isfinite(p[i])) return false; return true; }

// Symmetric positive-definite n x n (row-major, n <= 6) inverse by Cholesky; returns false if not PD.
bool spd_inverse(const double* A, int n, double* Ainv) {
    double L[36];
    for (int i = 0; i < n; ++i)
        for (int j = 0; j <= i; ++j) {
            double s = A[i * n + j];
            for (int k = 0; k < j; ++k) s -= L[i * n + k] * L[j * n + k];
            if (i == j) { if (!(s > 0.0)) return false; L[i * n + i] = std::sqrt(s); }
            else L[i * n + j] = s / L[j * n + j];
        }
    for (int c = 0; c < n; ++c) {  // solve L L^T x = e_c
        double y[6];
        for (int i = 0; i < n; ++i) { double s = (i == c) ? 1.0 : 0.0; for (int k = 0; k < i; ++k) s -= L[i * n + k] * y[k]; y[i] = s / L[i * n + i]; }
        for (int i = n - 1; i >= 0; --i) { double s = y[i]; for (int k = i + 1; k < n; ++k) s -= L[k * n + i] * Ainv[k * n + c]; Ainv[i * n + c] = s / L[i * n + i]; }
    }
    return true;
}

// In-place dense Cholesky A = L L^T (lower, row-major, leading dimension n), blocked right-looking, OpenMP over the
// trailing update.  Stands in for Eigen::SimplicialLDLT on the reduced camera matrix (BundleAdjuster.h:172-173):
// both are exact factorisations, so they agree to rounding.
bool dense_cholesky(double* A, int64_t n) {
    const int64_t B = 96;
    for (int64_t k0 = 0; k0 < n; k0 += B) {
        const int64_t kb = std::min(B, n - k0);
        for (int64_t j = k0; j < k0 + kb; ++j) {          // diagonal block, unblocked
            double d = A[j * n + j];
            for (int64_t k = k0; k < j; ++k) d -= A[j * n + k] * A[j * n + k];
            if (!(d > 0.0) || !std::isfinite(d)) return false;
            d = std::sqrt(d); A[j * n + j] = d;
            for (int64_t i = j + 1; i < k0 + kb; ++i) {
                double s = A[i * n + j];
                for (int64_t k = k0; k < j; ++k) s -= A[i * n + k] * A[j * n + k];
                A[i * n + j] = s / d;
            }
        }
        const int64_t r0 = k0 + kb;
#pragma omp parallel for schedule(static)
        for (int64_t i = r0; i < n; ++i)                   // panel: L21 = A21 L11^-T
            for (int64_t j = k0; j < k0 + kb; ++j) {
                double s = A[i * n + j];
                for (int64_t k = k0; k < j; ++k) s -= A[i * n + k] * A[j * n + k];
                A[i * n + j] = s / A[j * n + j];
            }
#pragma omp parallel for schedule(dynamic, 8)
        for (int64_t i = r0; i < n; ++i) {                 // trailing update (lower part): A22 -= L21 L21^T
            const double* li = &A[i * n + k0];
            for (int64_t j = r0; j <= i; ++j) {
                const double* lj = &A[j * n + k0];
                double s = 0.0;
#pragma omp simd reduction(+ : s)
                for (int64_t k = 0; k < kb; ++k) s += li[k] * lj[k];
                A[i * n + j] -= s;
            }
        }
    }
    return true;
}

void dense_cholesky_solve(const double* L, int64_t n, double* x) {  // x <- (L L^T)^-1 x
    for (int64_t i = 0; i < n; ++i) { double s = x[i]; for (int64_t k = 0; k < i; ++k) s -= L[i * n + k] * x[k]; x[i] = s / L[i * n + i]; }
    for (int64_t i = n - 1; i >= 0; --i) { double s = x[i]; for (int64_t k = i + 1; k < n; ++k) s -= L[k * n + i] * x[k]; x[i] = s / L[i * n + i]; }
}

}  // namespace

// ------------------------------------------------------------------------------------------------------------
struct ba_engine {
    ba_options opt;
    bool has_problem = false;
    int Nc = 0, Ng = 0, Np = 0, No = 0, ni = 4;   // ni: size of the intrinsics block of the model (4 pinhole, 3 Snavely)
    vector<double> pose0, intr0, pt0, pose, intr, pt;
    vector<int32_t> cam_group, obs_cam, obs_pt;
    vector<uint8_t> cam_fixed;
    vector<double> obs_uv;
    // index structure (canonical)
    vector<int32_t> pt_off, cam_off, perm_by_cam;
    vector<uint8_t> cam_free, grp_free, pt_free;   // free = not constant AND referenced by >= 1 residual (Ceres drops the rest)
    // evaluation buffers: robustified residuals and Jacobian blocks, J column-scaled once jacobi scaling is set
    vector<double> r, Jc, Jp, Ji;
    vector<double> sc_cam, sc_intr, sc_pt;          // Jacobi column scales (1 when off)
    vector<ba_iter_record> traj;
    ba_summary summary{};
    std::string err;
    int nthreads = 1;
    int nf() const { return 6 * Nc + (opt.intrinsics_free ? ni * Ng : 0); }
};

namespace {

int fail(ba_engine* e, int code, const std::string& msg) { if (e) e->err = msg; else g_create_error = msg; return code; }

void build_structure(ba_engine* e) {
    const int Nc = e->Nc, Np = e->Np, No = e->No;
    e->pt_off.assign(size_t(Np) + 1, 0);
    e->cam_off.assign(size_t(Nc) + 1, 0);
    for (int o = 0; o < No; ++o) { ++e->pt_off[e->obs_pt[o] + 1]; ++e->cam_off[e->obs_cam[o] + 1]; }
    for (int p = 0; p < Np; ++p) e->pt_off[p + 1] += e->pt_off[p];
    for (int c = 0; c < Nc; ++c) e->cam_off[c + 1] += e->cam_off[c];
    e->perm_by_cam.resize(No);
    vector<int32_t> cur(e->cam_off.begin(), e->cam_off.end() - 1);
    for (int o = 0; o < No; ++o) e->perm_by_cam[cur[e->obs_cam[o]]++] = o;   // stable: observation ids ascending within a camera
    e->cam_free.assign(Nc, 0); e->grp_free.assign(e->Ng, 0); e->pt_free.assign(Np, 0);
    for (int c = 0; c < Nc; ++c) e->cam_free[c] = (!e->cam_fixed[c] && e->cam_off[c + 1] > e->cam_off[c]) ? 1 : 0;
    for (int p = 0; p < Np; ++p) e->pt_free[p] = (e->pt_off[p + 1] > e->pt_off[p]) ? 1 : 0;
    if (e->opt.intrinsics_free)
        for (int c = 0; c < Nc; ++c) if (e->cam_off[c + 1] > e->cam_off[c]) e->grp_free[e->cam_group[c]] = 1;
}

// Evaluate all residual blocks at (pose, intr, pt): cost = 1/2 sum rho(||r||^2); optionally the corrected residuals
// and Jacobian blocks (Ceres' Corrector with rho'' <= 0: r <- sqrt(rho') r, J <- sqrt(rho') J — SURVEY a6).
// Columns of constant blocks are zero (Ceres removes them from the reduced program).  Returns false when any
// residual or Jacobian entry is non-finite (Ceres: evaluation failure).
template <class Cost>
bool evaluate_model(ba_engine* e, const double* pose, const double* intr, const double* pt, bool want_j, double* cost_out,
                    double* r, double* Jc, double* Jp, double* Ji, bool apply_scale) {
    const int No = e->No, ni = e->ni;
    const int kind = e->opt.robust_kind; const double a = e->opt.robust_a;
    const int nt = e->nthreads;
    vector<double> partial(nt, 0.0);
    vector<int> bad(nt, 0);
#pragma omp parallel for schedule(static) num_threads(nt)
    for (int o = 0; o < No; ++o) {
#ifdef _OPENMP
        const int tid = omp_get_thread_num();
#else
        const int tid = 0;
#endif
        const int c = e->obs_cam[o], p = e->obs_pt[o], g = e->cam_group[c];
        Cost f{e->obs_uv[2 * o], e->obs_uv[2 * o + 1], e->opt.taylor_rule};
        double res[2];
        double jac[2][13];
        if (want_j) {
            using J13 = oracle::Jet<13>;
            J13 ji[4], je[6], jx[3], jr[2];
            for (int k = 0; k < Cost::kIntr; ++k) ji[k] = J13(intr[4 * g + k], k);
            for (int k = 0; k < 6; ++k) je[k] = J13(pose[6 * c + k], 4 + k);
            for (int k = 0; k < 3; ++k) jx[k] = J13(pt[3 * p + k], 10 + k);
            f(ji, je, jx, jr);
            for (int q = 0; q < 2; ++q) { res[q] = jr[q].a; for (int k = 0; k < 13; ++k) jac[q][k] = jr[q].v[k]; }
            if (e->opt.lm_policy == BA_POLICY_G2O) {   // left twist (omega, upsilon): dP/d(omega) = -[P]x, dP/d(upsilon) = I  =>  row_omega = P x row_t
                double P[3];
                oracle::angle_axis_rotate_point(&pose[6 * c], &pt[3 * p], P, e->opt.taylor_rule);
                P[0] += pose[6 * c + 3]; P[1] += pose[6 * c + 4]; P[2] += pose[6 * c + 5];
                for (int q = 0; q < 2; ++q) { const double* a = &jac[q][7];
                    jac[q][4] = P[1] * a[2] - P[2] * a[1]; jac[q][5] = P[2] * a[0] - P[0] * a[2]; jac[q][6] = P[0] * a[1] - P[1] * a[0]; }
            }
            if (!finite_all(&jac[0][0], 26)) bad[tid] = 1;
        } else {
            f(&intr[4 * g], &pose[6 * c], &pt[3 * p], res);
        }
        if (!finite_all(res, 2)) { bad[tid] = 1; continue; }
        const double s = res[0] * res[0] + res[1] * res[1];
        double rho[3];
        oracle::loss_evaluate(kind, a, s, rho);
        partial[tid] += 0.5 * rho[0];
        if (!want_j && !r) continue;
        const double w = std::sqrt(rho[1]);
        if (r) { r[2 * o] = w * res[0]; r[2 * o + 1] = w * res[1]; }
        if (want_j) {
            const bool cf = e->cam_free[c], gf = e->opt.intrinsics_free && e->grp_free[g];
            for (int q = 0; q < 2; ++q) {
                for (int k = 0; k < ni; ++k) Ji[(2 * o + q) * ni + k] = gf ? w * jac[q][k] * (apply_scale ? e->sc_intr[ni * g + k] : 1.0) : 0.0;
                for (int k = 0; k < 6; ++k) Jc[(2 * o + q) * 6 + k] = cf ? w * jac[q][4 + k] * (apply_scale ? e->sc_cam[6 * c + k] : 1.0) : 0.0;
                for (int k = 0; k < 3; ++k) Jp[(2 * o + q) * 3 + k] = w * jac[q][10 + k] * (apply_scale ? e->sc_pt[3 * p + k] : 1.0);
            }
        }
    }
    double cost = 0.0; int anybad = 0;
    for (int t = 0; t < nt; ++t) { cost += partial[t]; anybad |= bad[t]; }
    *cost_out = cost;
    return !anybad && std::isfinite(cost);
}

bool evaluate(ba_engine* e, const double* pose, const double* intr, const double* pt, bool want_j, double* cost,
              double* r, double* Jc, double* Jp, double* Ji, bool apply_scale) {
    if (e->opt.model == BA_MODEL_SNAVELY)
        return evaluate_model<oracle::SnavelyCost>(e, pose, intr, pt, want_j, cost, r, Jc, Jp, Ji, apply_scale);
    return evaluate_model<oracle::PinholeCost>(e, pose, intr, pt, want_j, cost, r, Jc, Jp, Ji, apply_scale);
}

// g = J^T r over (pose | intrinsics | point); J may be column-scaled (then g is the scaled gradient).
void gradient(const ba_engine* e, const double* r, const double* Jc, const double* Jp, const double* Ji,
              double* g_cam, double* g_intr, double* g_pt) {
    const int ni = e->ni;
    std::fill(g_cam, g_cam + 6 * e->Nc, 0.0); std::fill(g_intr, g_intr + ni * e->Ng, 0.0); std::fill(g_pt, g_pt + 3 * e->Np, 0.0);
    for (int o = 0; o < e->No; ++o) {
        const int c = e->obs_cam[o], p = e->obs_pt[o], g = e->cam_group[c];
        for (int q = 0; q < 2; ++q) {
            const double rq = r[2 * o + q];
            for (int k = 0; k < 6; ++k) g_cam[6 * c + k] += Jc[(2 * o + q) * 6 + k] * rq;
            for (int k = 0; k < ni; ++k) g_intr[ni * g + k] += Ji[(2 * o + q) * ni + k] * rq;
            for (int k = 0; k < 3; ++k) g_pt[3 * p + k] += Jp[(2 * o + q) * 3 + k] * rq;
        }
    }
}

}  // namespace

namespace {

// ---- linear algebra of one LM step: solve (J^T J + D^2) y = J^T r by eliminating every point (SURVEY §8 a-LM) ----
struct StepOut { bool ok = false; int pcg_iters = 0; double pcg_res = 0.0; };

inline void mat_t_mul_2xk_2xm(const double* A, int k, const double* B, int m, double* C) {  // C(k x m) = A^T B, A is 2xk, B is 2xm
    for (int i = 0; i < k; ++i) for (int j = 0; j < m; ++j) C[i * m + j] = A[i] * B[j] + A[k + i] * B[m + j];
}

StepOut solve_step(ba_engine* e, double radius, const vector<double>& dg_cam, const vector<double>& dg_intr,
                   const vector<double>& dg_pt, vector<double>& y_cam, vector<double>& y_intr, vector<double>& y_pt) {
    StepOut out;
    const int Nc = e->Nc, Ng = e->Ng, Np = e->Np, ni = e->ni;
    const bool fi = e->opt.intrinsics_free != 0;
    const int nf = e->nf();
    const double* r = e->r.data(); const double* Jc = e->Jc.data(); const double* Jp = e->Jp.data(); const double* Ji = e->Ji.data();
    const int nt = e->nthreads;

    // 1. per point: Hpp = E^T E + Dp^2, its inverse, g_p = E^T r, v_p = Hpp^-1 g_p
    vector<double> Hinv(size_t(Np) * 9), gp(size_t(Np) * 3), vp(size_t(Np) * 3);
    int bad = 0;
#pragma omp parallel for schedule(static) num_threads(nt) reduction(| : bad)
    for (int p = 0; p < Np; ++p) {
        double H[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
        for (int o = e->pt_off[p]; o < e->pt_off[p + 1]; ++o)
            for (int q = 0; q < 2; ++q) {
                const double* jp = &Jp[(2 * o + q) * 3];
                for (int a = 0; a < 3; ++a) { g[a] += jp[a] * r[2 * o + q]; for (int b = 0; b < 3; ++b) H[a * 3 + b] += jp[a] * jp[b]; }
            }
        for (int a = 0; a < 3; ++a) H[a * 4] += dg_pt[3 * p + a] / radius;
        if (!spd_inverse(H, 3, &Hinv[size_t(p) * 9])) { bad |= 1; continue; }
        for (int a = 0; a < 3; ++a) {
            gp[3 * p + a] = g[a];
            vp[3 * p + a] = Hinv[size_t(p) * 9 + a * 3] * g[0] + Hinv[size_t(p) * 9 + a * 3 + 1] * g[1] + Hinv[size_t(p) * 9 + a * 3 + 2] * g[2];
        }
    }
    if (bad) return out;

    // damping of the f-blocks
    vector<double> Df(nf);
    for (int j = 0; j < 6 * Nc; ++j) Df[j] = dg_cam[j] / radius;
    if (fi) for (int j = 0; j < ni * Ng; ++j) Df[6 * Nc + j] = dg_intr[j] / radius;

    // reduced right-hand side b = F^T (r - E Hpp^-1 g_p)
    vector<double> b(nf, 0.0);
    for (int o = 0; o < e->No; ++o) {
        const int c = e->obs_cam[o], p = e->obs_pt[o], g = e->cam_group[c];
        for (int q = 0; q < 2; ++q) {
            const double* jp = &Jp[(2 * o + q) * 3];
            const double t = r[2 * o + q] - (jp[0] * vp[3 * p] + jp[1] * vp[3 * p + 1] + jp[2] * vp[3 * p + 2]);
            for (int k = 0; k < 6; ++k) b[6 * c + k] += Jc[(2 * o + q) * 6 + k] * t;
            if (fi) for (int k = 0; k < ni; ++k) b[6 * Nc + ni * g + k] += Ji[(2 * o + q) * ni + k] * t;
        }
    }

    vector<double> yf(nf, 0.0);
    if (e->opt.linear_solver != BA_LINEAR_PCG) {   // EXACT, and AUTO read as EXACT
        // 2a. explicit S = F^T F + Df^2 - sum_p (F^T E_p) Hpp^-1 (E_p^T F), dense; camera rows in parallel
        vector<double> S(size_t(nf) * nf, 0.0);
#pragma omp parallel for schedule(dynamic, 4) num_threads(nt)
        for (int c = 0; c < Nc; ++c) {
            double* Srow = &S[size_t(6 * c) * nf];
            for (int idx = e->cam_off[c]; idx < e->cam_off[c + 1]; ++idx) {
                const int o = e->perm_by_cam[idx], p = e->obs_pt[o], g = e->cam_group[c];
                const double* jc = &Jc[size_t(2 * o) * 6]; const double* jp = &Jp[size_t(2 * o) * 3]; const double* ji = &Ji[size_t(2 * o) * ni];
                double FtF[36]; mat_t_mul_2xk_2xm(jc, 6, jc, 6, FtF);
                for (int a = 0; a < 6; ++a) for (int k = 0; k < 6; ++k) Srow[size_t(a) * nf + 6 * c + k] += FtF[a * 6 + k];
                if (fi) { double FtI[24]; mat_t_mul_2xk_2xm(jc, 6, ji, ni, FtI);
                          for (int a = 0; a < 6; ++a) for (int k = 0; k < ni; ++k) Srow[size_t(a) * nf + 6 * Nc + ni * g + k] += FtI[a * ni + k]; }
                double T[18], TH[18]; mat_t_mul_2xk_2xm(jc, 6, jp, 3, T);
                const double* Hi = &Hinv[size_t(p) * 9];
                for (int a = 0; a < 6; ++a) for (int k = 0; k < 3; ++k) TH[a * 3 + k] = T[a * 3] * Hi[k] + T[a * 3 + 1] * Hi[3 + k] + T[a * 3 + 2] * Hi[6 + k];
                for (int o2 = e->pt_off[p]; o2 < e->pt_off[p + 1]; ++o2) {
                    const int c2 = e->obs_cam[o2], g2 = e->cam_group[c2];
                    double W[18]; mat_t_mul_2xk_2xm(&Jp[size_t(2 * o2) * 3], 3, &Jc[size_t(2 * o2) * 6], 6, W);  // 3x6
                    for (int a = 0; a < 6; ++a) for (int k = 0; k < 6; ++k)
                        Srow[size_t(a) * nf + 6 * c2 + k] -= TH[a * 3] * W[k] + TH[a * 3 + 1] * W[6 + k] + TH[a * 3 + 2] * W[12 + k];
                    if (fi) { double Wi[12]; mat_t_mul_2xk_2xm(&Jp[size_t(2 * o2) * 3], 3, &Ji[size_t(2 * o2) * ni], ni, Wi);  // 3 x ni
                              for (int a = 0; a < 6; ++a) for (int k = 0; k < ni; ++k)
                                  Srow[size_t(a) * nf + 6 * Nc + ni * g2 + k] -= TH[a * 3] * Wi[k] + TH[a * 3 + 1] * Wi[ni + k] + TH[a * 3 + 2] * Wi[2 * ni + k]; }
                }
            }
        }
        if (fi) {  // intrinsics rows: border by symmetry, corner block by a sweep over points
            for (int j = 0; j < ni * Ng; ++j) for (int k = 0; k < 6 * Nc; ++k) S[size_t(6 * Nc + j) * nf + k] = S[size_t(k) * nf + 6 * Nc + j];
            vector<double> U(size_t(Ng) * ni * 3);
            vector<int> touched;
            for (int p = 0; p < Np; ++p) {
                touched.clear();
                for (int o = e->pt_off[p]; o < e->pt_off[p + 1]; ++o) {
                    const int g = e->cam_group[e->obs_cam[o]];
                    if (std::find(touched.begin(), touched.end(), g) == touched.end()) { touched.push_back(g); std::fill(&U[size_t(g) * ni * 3], &U[size_t(g + 1) * ni * 3], 0.0); }
                    double II[16], IP[12];
                    mat_t_mul_2xk_2xm(&Ji[size_t(2 * o) * ni], ni, &Ji[size_t(2 * o) * ni], ni, II);
                    mat_t_mul_2xk_2xm(&Ji[size_t(2 * o) * ni], ni, &Jp[size_t(2 * o) * 3], 3, IP);
                    for (int a = 0; a < ni; ++a) for (int k = 0; k < ni; ++k) S[size_t(6 * Nc + ni * g + a) * nf + 6 * Nc + ni * g + k] += II[a * ni + k];
                    for (int a = 0; a < ni * 3; ++a) U[size_t(g) * ni * 3 + a] += IP[a];
                }
                const double* Hi = &Hinv[size_t(p) * 9];
                for (int g1 : touched) for (int g2 : touched)
                    for (int a = 0; a < ni; ++a) {
                        const double* u1 = &U[size_t(g1) * ni * 3 + a * 3];
                        const double uh[3] = {u1[0] * Hi[0] + u1[1] * Hi[3] + u1[2] * Hi[6], u1[0] * Hi[1] + u1[1] * Hi[4] + u1[2] * Hi[7], u1[0] * Hi[2] + u1[1] * Hi[5] + u1[2] * Hi[8]};
                        for (int k = 0; k < ni; ++k) { const double* u2 = &U[size_t(g2) * ni * 3 + k * 3];
                            S[size_t(6 * Nc + ni * g1 + a) * nf + 6 * Nc + ni * g2 + k] -= uh[0] * u2[0] + uh[1] * u2[1] + uh[2] * u2[2]; }
                    }
            }
        }
        for (int j = 0; j < nf; ++j) S[size_t(j) * nf + j] += Df[j];
        if (!dense_cholesky(S.data(), nf)) return out;
        yf = b;
        dense_cholesky_solve(S.data(), nf, yf.data());
    } else {
        // 2b. PCG on the implicit Schur complement (what the CUDA engine does; used here for shapes whose dense S does not fit)
        vector<double> Minv;  // block-diagonal preconditioner, 6x6 per pose (+ ni x ni per group)
        vector<double> blk(size_t(Nc) * 36, 0.0), blki(size_t(Ng) * ni * ni, 0.0);
        for (int o = 0; o < e->No; ++o) {
            const int c = e->obs_cam[o], g = e->cam_group[c];
            double A[36]; mat_t_mul_2xk_2xm(&Jc[size_t(2 * o) * 6], 6, &Jc[size_t(2 * o) * 6], 6, A);
            if (e->opt.preconditioner == BA_PRECOND_SCHUR_JACOBI) {   // diagonal block of S: subtract (F^T E) Hpp^-1 (E^T F) of same-camera pairs
                const int p = e->obs_pt[o]; const double* Hi = &Hinv[size_t(p) * 9];
                double T[18]; mat_t_mul_2xk_2xm(&Jc[size_t(2 * o) * 6], 6, &Jp[size_t(2 * o) * 3], 3, T);
                for (int o2 = e->pt_off[p]; o2 < e->pt_off[p + 1]; ++o2) {
                    if (e->obs_cam[o2] != c) continue;
                    double T2[18]; mat_t_mul_2xk_2xm(&Jc[size_t(2 * o2) * 6], 6, &Jp[size_t(2 * o2) * 3], 3, T2);
                    for (int a = 0; a < 6; ++a) { const double th[3] = {T[a * 3] * Hi[0] + T[a * 3 + 1] * Hi[3] + T[a * 3 + 2] * Hi[6], T[a * 3] * Hi[1] + T[a * 3 + 1] * Hi[4] + T[a * 3 + 2] * Hi[7], T[a * 3] * Hi[2] + T[a * 3 + 1] * Hi[5] + T[a * 3 + 2] * Hi[8]};
                        for (int k = 0; k < 6; ++k) A[a * 6 + k] -= th[0] * T2[k * 3] + th[1] * T2[k * 3 + 1] + th[2] * T2[k * 3 + 2]; }
                }
            }
            for (int a = 0; a < 36; ++a) blk[size_t(c) * 36 + a] += A[a];
            if (fi) { double II[16]; mat_t_mul_2xk_2xm(&Ji[size_t(2 * o) * ni], ni, &Ji[size_t(2 * o) * ni], ni, II); for (int a = 0; a < ni * ni; ++a) blki[size_t(g) * ni * ni + a] += II[a]; }
        }
        for (int c = 0; c < Nc; ++c) { for (int a = 0; a < 6; ++a) blk[size_t(c) * 36 + a * 7] += Df[6 * c + a];
            double inv[36]; if (!spd_inverse(&blk[size_t(c) * 36], 6, inv)) return out; std::memcpy(&blk[size_t(c) * 36], inv, sizeof inv); }
        if (fi) for (int g = 0; g < Ng; ++g) { for (int a = 0; a < ni; ++a) blki[size_t(g) * ni * ni + a * (ni + 1)] += Df[6 * Nc + ni * g + a];
            double inv[16]; if (!spd_inverse(&blki[size_t(g) * ni * ni], ni, inv)) return out; std::memcpy(&blki[size_t(g) * ni * ni], inv, ni * ni * sizeof(double)); }
        auto precond = [&](const vector<double>& x, vector<double>& z) {
            for (int c = 0; c < Nc; ++c) for (int a = 0; a < 6; ++a) { double s = 0; for (int k = 0; k < 6; ++k) s += blk[size_t(c) * 36 + a * 6 + k] * x[6 * c + k]; z[6 * c + a] = s; }
            if (fi) for (int g = 0; g < Ng; ++g) for (int a = 0; a < ni; ++a) { double s = 0; for (int k = 0; k < ni; ++k) s += blki[size_t(g) * ni * ni + a * ni + k] * x[6 * Nc + ni * g + k]; z[6 * Nc + ni * g + a] = s; }
        };
        vector<vector<double>> ytl(nt, vector<double>(nf));
        auto matvec = [&](const vector<double>& x, vector<double>& y) {   // y = S x, one sweep over points
#pragma omp parallel num_threads(nt)
            {
#ifdef _OPENMP
                const int tid = omp_get_thread_num();
#else
                const int tid = 0;
#endif
                vector<double>& yl = ytl[tid]; std::fill(yl.begin(), yl.end(), 0.0);
                vector<double> u;
#pragma omp for schedule(static)
                for (int p = 0; p < Np; ++p) {
                    const int o0 = e->pt_off[p], o1 = e->pt_off[p + 1];
                    u.assign(size_t(o1 - o0) * 2, 0.0);
                    double w[3] = {0, 0, 0};
                    for (int o = o0; o < o1; ++o) { const int c = e->obs_cam[o], g = e->cam_group[c];
                        for (int q = 0; q < 2; ++q) { double s = 0; for (int k = 0; k < 6; ++k) s += Jc[(2 * o + q) * 6 + k] * x[6 * c + k];
                            if (fi) for (int k = 0; k < ni; ++k) s += Ji[(2 * o + q) * ni + k] * x[6 * Nc + ni * g + k];
                            u[(o - o0) * 2 + q] = s; for (int a = 0; a < 3; ++a) w[a] += Jp[(2 * o + q) * 3 + a] * s; } }
                    const double* Hi = &Hinv[size_t(p) * 9];
                    const double v[3] = {Hi[0] * w[0] + Hi[1] * w[1] + Hi[2] * w[2], Hi[3] * w[0] + Hi[4] * w[1] + Hi[5] * w[2], Hi[6] * w[0] + Hi[7] * w[1] + Hi[8] * w[2]};
                    for (int o = o0; o < o1; ++o) { const int c = e->obs_cam[o], g = e->cam_group[c];
                        for (int q = 0; q < 2; ++q) { const double* jp = &Jp[(2 * o + q) * 3]; const double t = u[(o - o0) * 2 + q] - (jp[0] * v[0] + jp[1] * v[1] + jp[2] * v[2]);
                            for (int k = 0; k < 6; ++k) yl[6 * c + k] += Jc[(2 * o + q) * 6 + k] * t;
                            if (fi) for (int k = 0; k < ni; ++k) yl[6 * Nc + ni * g + k] += Ji[(2 * o + q) * ni + k] * t; } }
                }
            }
            for (int j = 0; j < nf; ++j) { double s = Df[j] * x[j]; for (int t = 0; t < nt; ++t) s += ytl[t][j]; y[j] = s; }
        };
        auto dotv = [&](const vector<double>& a, const vector<double>& c) { double s = 0; for (int j = 0; j < nf; ++j) s += a[j] * c[j]; return s; };
        vector<double> rr(b), z(nf), pdir(nf), q(nf);
        precond(rr, z); pdir = z;
        double rho = dotv(rr, z); const double rho0 = rho;
        double Q0 = 0.0; int k = 0; double rel = (rho0 > 0) ? 1.0 : 0.0;
        while (rho0 > 0 && k < e->opt.pcg_max_iterations) {
            matvec(pdir, q);
            const double pq = dotv(pdir, q);
            if (!(pq > 0) || !std::isfinite(pq)) break;
            const double alpha = rho / pq;
            for (int j = 0; j < nf; ++j) { yf[j] += alpha * pdir[j]; rr[j] -= alpha * q[j]; }
            precond(rr, z);
            const double rho_new = dotv(rr, z);
            ++k; rel = std::sqrt(std::max(0.0, rho_new) / rho0);
            if (e->opt.pcg_forcing) {   // Ceres' Q-criterion: Q = -1/2 x^T (b + r); stop when k (Q1 - Q0)/Q1 < eta
                double Q1 = 0; for (int j = 0; j < nf; ++j) Q1 += yf[j] * (b[j] + rr[j]); Q1 *= -0.5;
                const double zeta = k * (Q1 - Q0) / Q1; Q0 = Q1;
                if (zeta < e->opt.eta) break;
            }
            if (rel <= e->opt.pcg_rel_tolerance) break;
            const double beta = rho_new / rho; rho = rho_new;
            for (int j = 0; j < nf; ++j) pdir[j] = z[j] + beta * pdir[j];
        }
        out.pcg_iters = k; out.pcg_res = rel;
    }

    // 3. back-substitution: y_p = Hpp^-1 (g_p - E_p^T F y_f)
    for (int j = 0; j < 6 * Nc; ++j) y_cam[j] = yf[j];
    for (int j = 0; j < ni * Ng; ++j) y_intr[j] = fi ? yf[6 * Nc + j] : 0.0;
#pragma omp parallel for schedule(static) num_threads(nt)
    for (int p = 0; p < Np; ++p) {
        double w[3] = {gp[3 * p], gp[3 * p + 1], gp[3 * p + 2]};
        for (int o = e->pt_off[p]; o < e->pt_off[p + 1]; ++o) { const int c = e->obs_cam[o], g = e->cam_group[c];
            for (int q = 0; q < 2; ++q) { double s = 0; for (int k = 0; k < 6; ++k) s += Jc[(2 * o + q) * 6 + k] * y_cam[6 * c + k];
                if (fi) for (int k = 0; k < ni; ++k) s += Ji[(2 * o + q) * ni + k] * y_intr[ni * g + k];
                for (int a = 0; a < 3; ++a) w[a] -= Jp[(2 * o + q) * 3 + a] * s; } }
        const double* Hi = &Hinv[size_t(p) * 9];
        for (int a = 0; a < 3; ++a) y_pt[3 * p + a] = Hi[a * 3] * w[0] + Hi[a * 3 + 1] * w[1] + Hi[a * 3 + 2] * w[2];
    }
    out.ok = finite_all(yf.data(), nf) && finite_all(y_pt.data(), 3 * Np);
    return out;
}

}  // namespace

namespace {

// ---- Ceres 2.2 TrustRegionMinimizer + LevenbergMarquardtStrategy, restated (SURVEY §8 a-LM; every rule there is flagged
//      "from upstream knowledge": the source is not under /root/reference) ----
struct FreeNorm {
    const ba_engine* e;
    template <class F> double norm(F f) const {   // sqrt(sum over free scalars of f(kind, index)^2), order: points, poses, intrinsics
        double s = 0.0;
        for (int p = 0; p < e->Np; ++p) if (e->pt_free[p]) for (int a = 0; a < 3; ++a) { const double v = f(2, 3 * p + a); s += v * v; }
        for (int c = 0; c < e->Nc; ++c) if (e->cam_free[c]) for (int a = 0; a < 6; ++a) { const double v = f(0, 6 * c + a); s += v * v; }
        for (int g = 0; g < e->Ng; ++g) if (e->grp_free[g]) for (int a = 0; a < e->ni; ++a) { const double v = f(1, 4 * g + a); s += v * v; }
        return std::sqrt(s);
    }
};

void column_sqnorms(const ba_engine* e, vector<double>& n_cam, vector<double>& n_intr, vector<double>& n_pt) {
    const int ni = e->ni;
    n_cam.assign(size_t(6) * e->Nc, 0.0); n_intr.assign(size_t(ni) * e->Ng, 0.0); n_pt.assign(size_t(3) * e->Np, 0.0);
    for (int o = 0; o < e->No; ++o) {
        const int c = e->obs_cam[o], p = e->obs_pt[o], g = e->cam_group[c];
        for (int q = 0; q < 2; ++q) {
            for (int k = 0; k < 6; ++k) { const double v = e->Jc[(2 * o + q) * 6 + k]; n_cam[6 * c + k] += v * v; }
            for (int k = 0; k < ni; ++k) { const double v = e->Ji[(2 * o + q) * ni + k]; n_intr[ni * g + k] += v * v; }
            for (int k = 0; k < 3; ++k) { const double v = e->Jp[(2 * o + q) * 3 + k]; n_pt[3 * p + k] += v * v; }
        }
    }
}

// ---- g2o policy: OptimizationAlgorithmLevenberg + VertexSE3Expmap, restated from g2o's published algorithm ----
void so3_exp(const double w[3], double R[9]) {
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = std::sqrt(th2);
    double a, b;   // R = I + a [w]x + b [w]x^2
    if (th < 1e-5) { a = 1.0 - th2 / 6.0; b = 0.5 - th2 / 24.0; } else { a = std::sin(th) / th; b = (1.0 - std::cos(th)) / th2; }
    const double K[9] = {0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0};
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) {
        double k2 = 0; for (int m = 0; m < 3; ++m) k2 += K[i * 3 + m] * K[m * 3 + j];
        R[i * 3 + j] = (i == j ? 1.0 : 0.0) + a * K[i * 3 + j] + b * k2;
    }
}
void so3_log(const double R[9], double w[3]) {   // through the unit quaternion, angle in [0, pi]
    double q[4]; const double tr = R[0] + R[4] + R[8];
    if (tr > 0.0) { double t = std::sqrt(tr + 1.0); q[0] = 0.5 * t; t = 0.5 / t; q[1] = (R[7] - R[5]) * t; q[2] = (R[2] - R[6]) * t; q[3] = (R[3] - R[1]) * t; }
    else { int i = 0; if (R[4] > R[0]) i = 1; if (R[8] > R[4 * i]) i = 2; const int j = (i + 1) % 3, k = (j + 1) % 3;
           double t = std::sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0); q[1 + i] = 0.5 * t; t = 0.5 / t;
           q[0] = (R[3 * k + j] - R[3 * j + k]) * t; q[1 + j] = (R[3 * j + i] + R[3 * i + j]) * t; q[1 + k] = (R[3 * k + i] + R[3 * i + k]) * t; }
    if (q[0] < 0.0) for (double& v : q) v = -v;
    const double n = std::sqrt(q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    if (n == 0.0) { w[0] = w[1] = w[2] = 0.0; return; }
    const double ang = 2.0 * std::atan2(n, q[0]);
    for (int i = 0; i < 3; ++i) w[i] = ang * q[1 + i] / n;
}
// pose <- exp(omega, upsilon) * pose  (VertexSE3Expmap::oplus): R' = exp(w) R, t' = exp(w) t + V(w) u
void se3_left_update(const double pose[6], const double w[3], const double u[3], double out[6]) {
    double Rd[9], R[9], Rn[9];
    so3_exp(w, Rd); so3_exp(pose, R);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { double v = 0; for (int m = 0; m < 3; ++m) v += Rd[i * 3 + m] * R[m * 3 + j]; Rn[i * 3 + j] = v; }
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = std::sqrt(th2);
    double b, c;   // V = I + b [w]x + c [w]x^2
    if (th < 1e-5) { b = 0.5 - th2 / 24.0; c = 1.0 / 6.0 - th2 / 120.0; } else { b = (1.0 - std::cos(th)) / th2; c = (th - std::sin(th)) / (th2 * th); }
    const double wu[3] = {w[1] * u[2] - w[2] * u[1], w[2] * u[0] - w[0] * u[2], w[0] * u[1] - w[1] * u[0]};
    const double wwu[3] = {w[1] * wu[2] - w[2] * wu[1], w[2] * wu[0] - w[0] * wu[2], w[0] * wu[1] - w[1] * wu[0]};
    so3_log(Rn, out);
    for (int i = 0; i < 3; ++i) out[3 + i] = Rd[i * 3] * pose[3] + Rd[i * 3 + 1] * pose[4] + Rd[i * 3 + 2] * pose[5] + u[i] + b * wu[i] + c * wwu[i];
}

int lm_solve_g2o(ba_engine* e) {
    const ba_options& O = e->opt;
    const int Nc = e->Nc, Ng = e->Ng, Np = e->Np, No = e->No, ni = e->ni;
    const auto t0 = std::chrono::steady_clock::now();
    ba_summary& S = e->summary; std::memset(&S, 0, sizeof S);
    S.num_residuals = 2 * No;
    e->traj.clear();
    e->r.assign(size_t(2) * No, 0.0); e->Jc.assign(size_t(12) * No, 0.0); e->Jp.assign(size_t(6) * No, 0.0); e->Ji.assign(size_t(2) * ni * No, 0.0);
    e->sc_cam.assign(size_t(6) * Nc, 1.0); e->sc_intr.assign(size_t(ni) * Ng, 1.0); e->sc_pt.assign(size_t(3) * Np, 1.0);
    vector<double>&pose = e->pose, &pt = e->pt;
    vector<double> cpose(pose), cpt(pt), g_cam(6 * Nc), g_intr(ni * Ng), g_pt(3 * Np), y_cam(6 * Nc), y_intr(ni * Ng), y_pt(3 * Np);
    auto finish = [&](int term, const char* msg, double cost) {
        S.termination = term; S.usable = (term != BA_TERM_FAILURE); S.final_cost = cost; S.num_iterations = int(e->traj.size());
        std::snprintf(S.message, sizeof S.message, "%s", msg);
        S.solve_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        return BA_OK;
    };
    double cost = 0.0;
    if (!evaluate(e, pose.data(), e->intr.data(), pt.data(), true, &cost, e->r.data(), e->Jc.data(), e->Jp.data(), e->Ji.data(), false)) {
        S.initial_cost = cost;
        return finish(BA_TERM_FAILURE, "initial residual/Jacobian evaluation failed (non-finite)", cost);
    }
    S.initial_cost = cost;
    { ba_iter_record R{}; R.step_is_valid = 1; R.step_is_successful = 1; R.cost = cost; e->traj.push_back(R); S.num_successful_steps = 1; }
    double lambda = 0.0, ni_fac = 2.0;
    const vector<double> ones_c(6 * Nc, 1.0), ones_i(ni * Ng, 1.0), ones_p(3 * Np, 1.0);
    int record = 0;
    for (int iteration = 0; iteration < O.max_num_iterations; ++iteration) {
        if (O.target_chi2 > 0.0 && 2.0 * cost <= O.target_chi2) return finish(BA_TERM_TARGET_REACHED, "target chi2 reached", cost);
        gradient(e, e->r.data(), e->Jc.data(), e->Jp.data(), e->Ji.data(), g_cam.data(), g_intr.data(), g_pt.data());
        if (iteration == 0) {   // computeLambdaInit: tau * max diagonal of H
            vector<double> n_cam, n_intr, n_pt; column_sqnorms(e, n_cam, n_intr, n_pt);
            double mx = 0.0;
            for (double v : n_cam) mx = std::max(mx, v);
            for (double v : n_pt) mx = std::max(mx, v);
            lambda = 1e-5 * mx; ni_fac = 2.0;
        }
        double rho = 0.0; int qmax = 0;
        do {
            ba_iter_record R{}; R.iteration = ++record; R.radius_used = 1.0 / lambda; R.cost = cost;
            StepOut so = solve_step(e, 1.0 / lambda, ones_c, ones_i, ones_p, y_cam, y_intr, y_pt);
            R.pcg_iterations = so.pcg_iters; R.pcg_rel_residual = so.pcg_res; S.total_pcg_iterations += so.pcg_iters;
            double tempChi = std::numeric_limits<double>::max(), scale = 0.0;
            if (so.ok) {
                for (int c = 0; c < Nc; ++c) {
                    if (!e->cam_free[c]) { for (int k = 0; k < 6; ++k) cpose[6 * c + k] = pose[6 * c + k]; continue; }
                    const double w[3] = {-y_cam[6 * c], -y_cam[6 * c + 1], -y_cam[6 * c + 2]}, u[3] = {-y_cam[6 * c + 3], -y_cam[6 * c + 4], -y_cam[6 * c + 5]};
                    se3_left_update(&pose[6 * c], w, u, &cpose[6 * c]);
                }
                for (int j = 0; j < 3 * Np; ++j) cpt[j] = pt[j] - y_pt[j];
                double ccost = 0.0;
                if (evaluate(e, cpose.data(), e->intr.data(), cpt.data(), false, &ccost, nullptr, nullptr, nullptr, nullptr, false)) tempChi = 2.0 * ccost;
                for (int j = 0; j < 6 * Nc; ++j) scale += y_cam[j] * (lambda * y_cam[j] + g_cam[j]);   // dx.(lambda dx + b), dx = -y, b = -g
                for (int j = 0; j < 3 * Np; ++j) scale += y_pt[j] * (lambda * y_pt[j] + g_pt[j]);
            }
            R.step_is_valid = so.ok; R.candidate_cost = tempChi >= std::numeric_limits<double>::max() ? tempChi : 0.5 * tempChi;
            R.model_cost_change = 0.5 * scale;
            rho = (2.0 * cost - tempChi) / (scale + 1e-3);
            R.relative_decrease = rho;
            if (rho > 0.0 && std::isfinite(tempChi)) {
                const double alpha = std::min(1.0 - std::pow(2.0 * rho - 1.0, 3), 2.0 / 3.0);
                lambda *= std::max(1.0 / 3.0, alpha); ni_fac = 2.0;
                pose = cpose; pt = cpt;
                if (!evaluate(e, pose.data(), e->intr.data(), pt.data(), true, &cost, e->r.data(), e->Jc.data(), e->Jp.data(), e->Ji.data(), false))
                    return finish(BA_TERM_FAILURE, "residual/Jacobian evaluation failed after an accepted step", cost);
                R.step_is_successful = 1; R.cost = cost; ++S.num_successful_steps;
            } else {
                lambda *= ni_fac; ni_fac *= 2.0; ++S.num_unsuccessful_steps;
            }
            R.radius = 1.0 / lambda;
            e->traj.push_back(R);
            if (!std::isfinite(lambda)) break;
            ++qmax;
        } while (rho < 0.0 && qmax < 10);
        if (qmax == 10 || rho == 0.0 || !std::isfinite(lambda)) return finish(BA_TERM_CONVERGENCE, "g2o: no further improvement (trial limit, zero gain or lambda overflow)", cost);
    }
    return finish(BA_TERM_NO_CONVERGENCE, "maximum number of iterations reached", cost);
}

int lm_solve(ba_engine* e) {
    const ba_options& O = e->opt;
    const int Nc = e->Nc, Ng = e->Ng, Np = e->Np, No = e->No, ni = e->ni;
    const auto t0 = std::chrono::steady_clock::now();
    ba_summary& S = e->summary; std::memset(&S, 0, sizeof S);
    S.num_residuals = 2 * No;
    e->traj.clear();
    e->r.assign(size_t(2) * No, 0.0); e->Jc.assign(size_t(12) * No, 0.0); e->Jp.assign(size_t(6) * No, 0.0); e->Ji.assign(size_t(2) * ni * No, 0.0);
    e->sc_cam.assign(size_t(6) * Nc, 1.0); e->sc_intr.assign(size_t(ni) * Ng, 1.0); e->sc_pt.assign(size_t(3) * Np, 1.0);
    vector<double>&pose = e->pose, &intr = e->intr, &pt = e->pt;
    vector<double> cpose(pose), cintr(intr), cpt(pt);
    vector<double> g_cam(6 * Nc), g_intr(ni * Ng), g_pt(3 * Np);
    vector<double> dg_cam, dg_intr, dg_pt, y_cam(6 * Nc), y_intr(ni * Ng), y_pt(3 * Np);
    FreeNorm FN{e};
    auto finish = [&](int term, const char* msg, double cost) {
        S.termination = term; S.usable = (term != BA_TERM_FAILURE); S.final_cost = cost; S.num_iterations = int(e->traj.size());
        std::snprintf(S.message, sizeof S.message, "%s", msg);
        S.solve_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        return BA_OK;
    };
    // gradient max-norm exactly as the minimizer forms it: ||x - Plus(x, -g)||_inf with Plus(x,d) = x + d
    auto grad_max = [&]() {
        gradient(e, e->r.data(), e->Jc.data(), e->Jp.data(), e->Ji.data(), g_cam.data(), g_intr.data(), g_pt.data());
        double m = 0.0;
        auto upd = [&](double x, double gs, double sc) { const double g = gs / sc; m = std::max(m, std::fabs(x - (x + (-g)))); };
        for (int p = 0; p < Np; ++p) if (e->pt_free[p]) for (int a = 0; a < 3; ++a) upd(pt[3 * p + a], g_pt[3 * p + a], e->sc_pt[3 * p + a]);
        for (int c = 0; c < Nc; ++c) if (e->cam_free[c]) for (int a = 0; a < 6; ++a) upd(pose[6 * c + a], g_cam[6 * c + a], e->sc_cam[6 * c + a]);
        for (int g = 0; g < Ng; ++g) if (e->grp_free[g]) for (int a = 0; a < ni; ++a) upd(intr[4 * g + a], g_intr[ni * g + a], e->sc_intr[ni * g + a]);
        return m;
    };

    // ---- iteration 0 ----
    double cost = 0.0;
    if (!evaluate(e, pose.data(), intr.data(), pt.data(), true, &cost, e->r.data(), e->Jc.data(), e->Jp.data(), e->Ji.data(), false)) {
        S.initial_cost = cost;
        return finish(BA_TERM_FAILURE, "initial residual/Jacobian evaluation failed (non-finite)", cost);
    }
    S.initial_cost = cost;
    double gmax = grad_max();
    if (O.jacobi_scaling) {   // s_j = 1 / (1 + sqrt(sum_i J_ij^2)), frozen for the whole solve
        vector<double> n_cam, n_intr, n_pt; column_sqnorms(e, n_cam, n_intr, n_pt);
        for (size_t j = 0; j < n_cam.size(); ++j) e->sc_cam[j] = 1.0 / (1.0 + std::sqrt(n_cam[j]));
        for (size_t j = 0; j < n_intr.size(); ++j) e->sc_intr[j] = 1.0 / (1.0 + std::sqrt(n_intr[j]));
        for (size_t j = 0; j < n_pt.size(); ++j) e->sc_pt[j] = 1.0 / (1.0 + std::sqrt(n_pt[j]));
        for (int o = 0; o < No; ++o) { const int c = e->obs_cam[o], p = e->obs_pt[o], g = e->cam_group[c];
            for (int q = 0; q < 2; ++q) { for (int k = 0; k < 6; ++k) e->Jc[(2 * o + q) * 6 + k] *= e->sc_cam[6 * c + k];
                for (int k = 0; k < ni; ++k) e->Ji[(2 * o + q) * ni + k] *= e->sc_intr[ni * g + k];
                for (int k = 0; k < 3; ++k) e->Jp[(2 * o + q) * 3 + k] *= e->sc_pt[3 * p + k]; } }
    }
    double x_norm = FN.norm([&](int kind, int i) { return kind == 0 ? pose[i] : kind == 1 ? intr[i] : pt[i]; });
    double radius = O.initial_trust_region_radius, decrease_factor = 2.0;
    bool reuse_diagonal = false; int num_invalid = 0;
    {
        ba_iter_record R{}; R.iteration = 0; R.step_is_valid = 1; R.step_is_successful = 1; R.cost = cost; R.radius = radius; R.radius_used = radius; R.gradient_max_norm = gmax;
        e->traj.push_back(R); S.num_successful_steps = 1;
    }
    int iteration = 0;
    for (;;) {
        const ba_iter_record& last = e->traj.back();
        if (iteration >= O.max_num_iterations) return finish(BA_TERM_NO_CONVERGENCE, "maximum number of iterations reached", cost);
        if (last.step_is_successful && gmax <= O.gradient_tolerance) return finish(BA_TERM_CONVERGENCE, "gradient tolerance reached", cost);
        if (radius <= O.min_trust_region_radius) return finish(BA_TERM_CONVERGENCE, "minimum trust region radius reached", cost);
        if (O.target_chi2 > 0.0 && 2.0 * cost <= O.target_chi2) return finish(BA_TERM_TARGET_REACHED, "target chi2 reached", cost);
        ++iteration;
        ba_iter_record R{}; R.iteration = iteration; R.radius_used = radius; R.gradient_max_norm = gmax; R.cost = cost;

        // LevenbergMarquardtStrategy::ComputeStep
        if (!reuse_diagonal) {
            column_sqnorms(e, dg_cam, dg_intr, dg_pt);
            for (auto* v : {&dg_cam, &dg_intr, &dg_pt}) for (double& d : *v) d = std::min(std::max(d, O.min_lm_diagonal), O.max_lm_diagonal);
        }
        StepOut so = solve_step(e, radius, dg_cam, dg_intr, dg_pt, y_cam, y_intr, y_pt);
        reuse_diagonal = true;
        R.pcg_iterations = so.pcg_iters; R.pcg_rel_residual = so.pcg_res; S.total_pcg_iterations += so.pcg_iters;
        double mcc = 0.0;
        if (so.ok) {   // model_cost_change = -(J d)^T (r + J d / 2), d = -y (scaled space)
            double acc = 0.0;
            for (int o = 0; o < No; ++o) { const int c = e->obs_cam[o], p = e->obs_pt[o], g = e->cam_group[c];
                for (int q = 0; q < 2; ++q) { double m = 0.0;
                    for (int k = 0; k < 6; ++k) m -= e->Jc[(2 * o + q) * 6 + k] * y_cam[6 * c + k];
                    for (int k = 0; k < ni; ++k) m -= e->Ji[(2 * o + q) * ni + k] * y_intr[ni * g + k];
                    for (int k = 0; k < 3; ++k) m -= e->Jp[(2 * o + q) * 3 + k] * y_pt[3 * p + k];
                    acc += m * (e->r[2 * o + q] + 0.5 * m); } }
            mcc = -acc;
        }
        R.model_cost_change = mcc;
        const bool valid = so.ok && mcc > 0.0;
        R.step_is_valid = valid;
        if (!valid) {   // HandleInvalidStep
            if (++num_invalid >= O.max_num_consecutive_invalid_steps) { e->traj.push_back(R); ++S.num_unsuccessful_steps; return finish(BA_TERM_FAILURE, "too many consecutive invalid steps", cost); }
            radius /= decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;
            R.radius = radius; e->traj.push_back(R); ++S.num_unsuccessful_steps;
            continue;
        }
        num_invalid = 0;
        // candidate x+ = x + s * delta_s
        for (int j = 0; j < 6 * Nc; ++j) cpose[j] = pose[j] - e->sc_cam[j] * y_cam[j];
        for (int g = 0; g < Ng; ++g) for (int a = 0; a < 4; ++a) cintr[4 * g + a] = intr[4 * g + a] - (a < ni ? e->sc_intr[ni * g + a] * y_intr[ni * g + a] : 0.0);
        for (int j = 0; j < 3 * Np; ++j) cpt[j] = pt[j] - e->sc_pt[j] * y_pt[j];
        double ccost = 0.0;
        if (!evaluate(e, cpose.data(), cintr.data(), cpt.data(), false, &ccost, nullptr, nullptr, nullptr, nullptr, false)) ccost = std::numeric_limits<double>::max();
        R.candidate_cost = ccost;
        R.step_norm = FN.norm([&](int kind, int i) { return kind == 0 ? pose[i] - cpose[i] : kind == 1 ? intr[i] - cintr[i] : pt[i] - cpt[i]; });
        // TrustRegionMinimizer::ParameterToleranceReached / FunctionToleranceReached, in this order, before the accept test, on every valid
        // step: Ceres 2.2 has NO "at least one successful step" guard on these two tests (SURVEY 8 a-LM lists one; that guard is not in the
        // 2.2 minimizer and is deliberately not implemented here) — restated from the published source, not pinned by a Ceres run.
        if (R.step_norm <= O.parameter_tolerance * (x_norm + O.parameter_tolerance)) return finish(BA_TERM_CONVERGENCE, "parameter tolerance reached", cost);
        if (std::fabs(cost - ccost) <= O.function_tolerance * cost) return finish(BA_TERM_CONVERGENCE, "function tolerance reached", cost);
        R.relative_decrease = (ccost >= std::numeric_limits<double>::max()) ? std::numeric_limits<double>::lowest() : (cost - ccost) / mcc;
        if (R.relative_decrease > O.min_relative_decrease) {   // HandleSuccessfulStep
            pose = cpose; intr = cintr; pt = cpt;
            x_norm = FN.norm([&](int kind, int i) { return kind == 0 ? pose[i] : kind == 1 ? intr[i] : pt[i]; });
            if (!evaluate(e, pose.data(), intr.data(), pt.data(), true, &cost, e->r.data(), e->Jc.data(), e->Jp.data(), e->Ji.data(), true))
                return finish(BA_TERM_FAILURE, "residual/Jacobian evaluation failed after an accepted step", cost);
            gmax = grad_max();
            radius = std::min(O.max_trust_region_radius, radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * R.relative_decrease - 1.0, 3)));
            decrease_factor = 2.0; reuse_diagonal = false;
            R.step_is_successful = 1; R.cost = cost; R.gradient_max_norm = gmax; ++S.num_successful_steps;
        } else {
            radius /= decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true; ++S.num_unsuccessful_steps;
        }
        R.radius = radius;
        e->traj.push_back(R);
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------------------
// C-ABI
extern "C" {

int ba_options_default(ba_options* o) {
    if (!o) return BA_ERR_INVALID_ARGUMENT;
    std::memset(o, 0, sizeof *o);
    o->abi_version = BA_ABI_VERSION;
    o->model = BA_MODEL_PINHOLE;
    o->robust_kind = BA_ROBUST_HUBER; o->robust_a = 4.0;          // BundleAdjuster.h:109
    o->intrinsics_free = 1;                                        // :106-107
    o->taylor_rule = BA_TAYLOR_THETA2_LE_EPS;
    o->max_num_iterations = 50; o->initial_trust_region_radius = 1e4; o->max_trust_region_radius = 1e16; o->min_trust_region_radius = 1e-32;
    o->min_relative_decrease = 1e-3; o->min_lm_diagonal = 1e-6; o->max_lm_diagonal = 1e32; o->max_num_consecutive_invalid_steps = 5;
    o->function_tolerance = 1e-6; o->gradient_tolerance = 1e-10; o->parameter_tolerance = 1e-8; o->jacobi_scaling = 1;
    o->linear_solver = BA_LINEAR_EXACT;                            // :172-173 (SPARSE_SCHUR, exact factorisation)
    o->preconditioner = BA_PRECOND_JACOBI;                         // :171
    o->pcg_max_iterations = 2000; o->pcg_rel_tolerance = 1e-12; o->pcg_forcing = 0; o->eta = 0.1;
    o->target_chi2 = 0.0; o->device = 0; o->num_threads = 1;       // :170
    o->rank = 0; o->world_size = 1; o->nccl_unique_id = nullptr; o->stream = nullptr; o->scatter_mode = 0; o->p2p_allreduce = 0; o->lm_policy = BA_POLICY_CERES;
    return BA_OK;
}

int ba_create(const ba_options* opt, ba_engine** out) {
    if (!opt || !out) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "null argument");
    if (opt->abi_version != BA_ABI_VERSION) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "abi_version mismatch");
    if (opt->world_size != 1) return fail(nullptr, BA_ERR_UNSUPPORTED, "the CPU oracle is single-process");
    if (opt->lm_policy == BA_POLICY_G2O && (opt->intrinsics_free || opt->model != BA_MODEL_PINHOLE)) return fail(nullptr, BA_ERR_UNSUPPORTED, "g2o policy: EdgeSE3ProjectXYZ has constant pinhole intrinsics (set intrinsics_free = 0)");
    ba_engine* e = new (std::nothrow) ba_engine();
    if (!e) return fail(nullptr, BA_ERR_INTERNAL, "out of memory");
    e->opt = *opt;
    e->ni = (opt->model == BA_MODEL_SNAVELY) ? 3 : 4;
    e->nthreads = std::max(1, opt->num_threads);
#ifndef _OPENMP
    e->nthreads = 1;
#endif
    *out = e;
    return BA_OK;
}

int ba_set_problem(ba_engine* e, const ba_problem* p) {
    if (!e || !p) return fail(e, BA_ERR_INVALID_ARGUMENT, "null argument");
    if (p->num_cameras < 0 || p->num_groups < 0 || p->num_points < 0 || p->num_observations < 0) return fail(e, BA_ERR_INVALID_ARGUMENT, "negative size");
    const int Nc = p->num_cameras, Ng = p->num_groups, Np = p->num_points, No = p->num_observations;
    if ((Nc && (!p->pose || !p->cam_group || !p->cam_fixed)) || (Ng && !p->intr) || (Np && !p->pt) || (No && (!p->obs_cam || !p->obs_pt || !p->obs_uv)))
        return fail(e, BA_ERR_INVALID_ARGUMENT, "null array");
    for (int c = 0; c < Nc; ++c) if (p->cam_group[c] < 0 || p->cam_group[c] >= Ng) return fail(e, BA_ERR_INVALID_ARGUMENT, "cam_group out of range");
    for (int o = 0; o < No; ++o) {
        if (p->obs_cam[o] < 0 || p->obs_cam[o] >= Nc || p->obs_pt[o] < 0 || p->obs_pt[o] >= Np) return fail(e, BA_ERR_INVALID_ARGUMENT, "observation index out of range");
        if (o && p->obs_pt[o] < p->obs_pt[o - 1]) return fail(e, BA_ERR_INVALID_ARGUMENT, "observations must be sorted by point id");
    }
    e->Nc = Nc; e->Ng = Ng; e->Np = Np; e->No = No;
    e->pose0.assign(p->pose, p->pose + size_t(6) * Nc); e->intr0.assign(p->intr, p->intr + size_t(4) * Ng); e->pt0.assign(p->pt, p->pt + size_t(3) * Np);
    e->cam_group.assign(p->cam_group, p->cam_group + Nc); e->cam_fixed.assign(p->cam_fixed, p->cam_fixed + Nc);
    e->obs_cam.assign(p->obs_cam, p->obs_cam + No); e->obs_pt.assign(p->obs_pt, p->obs_pt + No); e->obs_uv.assign(p->obs_uv, p->obs_uv + size_t(2) * No);
    e->pose = e->pose0; e->intr = e->intr0; e->pt = e->pt0;
    build_structure(e);
    e->traj.clear(); e->has_problem = true;
    return BA_OK;
}

int ba_append(ba_engine* e, const ba_delta* d) {
    if (!e || !d) return fail(e, BA_ERR_INVALID_ARGUMENT, "null argument");
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    ba_merge::Merged M;
    const std::string why = ba_merge::merge(e->Nc, e->Ng, e->Np, e->No, e->cam_group.data(), e->cam_fixed.data(), e->obs_cam.data(), e->obs_pt.data(), e->obs_uv.data(), *d, M);
    if (!why.empty()) return fail(e, BA_ERR_INVALID_ARGUMENT, "ba_append: " + why);
    vector<double> pose = e->pose, intr = e->intr, pt = e->pt;          // the current (optimised) state is kept
    pose.insert(pose.end(), d->pose, d->pose + size_t(6) * d->num_new_cameras);
    intr.insert(intr.end(), d->intr, d->intr + size_t(4) * d->num_new_groups);
    pt.insert(pt.end(), d->pt, d->pt + size_t(3) * d->num_new_points);
    ba_problem P{};
    P.num_cameras = M.Nc; P.num_groups = M.Ng; P.num_points = M.Np; P.num_observations = M.No;
    P.pose = pose.data(); P.intr = intr.data(); P.cam_group = M.cam_group.data(); P.cam_fixed = M.cam_fixed.data(); P.pt = pt.data();
    P.obs_cam = M.obs_cam.data(); P.obs_pt = M.obs_pt.data(); P.obs_uv = M.obs_uv.data();
    return ba_set_problem(e, &P);
}

int ba_update_parameters(ba_engine* e, int32_t kind, int32_t first, int32_t count, const double* values) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    if (kind < 0 || kind > 2) return fail(e, BA_ERR_INVALID_ARGUMENT, "ba_update_parameters: kind must be 0 (poses), 1 (intrinsics) or 2 (points)");
    const int w = kind == 0 ? 6 : (kind == 1 ? 4 : 3), n = kind == 0 ? e->Nc : (kind == 1 ? e->Ng : e->Np);
    if (first < 0 || count < 0 || int64_t(first) + count > n || (count && !values)) return fail(e, BA_ERR_INVALID_ARGUMENT, "ba_update_parameters: range out of bounds");
    vector<double>& cur = kind == 0 ? e->pose : (kind == 1 ? e->intr : e->pt);
    vector<double>& ini = kind == 0 ? e->pose0 : (kind == 1 ? e->intr0 : e->pt0);
    std::copy(values, values + size_t(w) * count, cur.begin() + size_t(w) * first);
    std::copy(values, values + size_t(w) * count, ini.begin() + size_t(w) * first);
    return BA_OK;
}

int ba_get_observations(ba_engine* e, int32_t* obs_cam, int32_t* obs_pt, double* obs_uv, int32_t cap, int32_t* n) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    if (n) *n = e->No;
    const int m = std::max(0, std::min<int>(cap, e->No));
    if (obs_cam) std::copy(e->obs_cam.begin(), e->obs_cam.begin() + m, obs_cam);
    if (obs_pt) std::copy(e->obs_pt.begin(), e->obs_pt.begin() + m, obs_pt);
    if (obs_uv) std::copy(e->obs_uv.begin(), e->obs_uv.begin() + size_t(2) * m, obs_uv);
    return BA_OK;
}

int ba_reset_state(ba_engine* e) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    e->pose = e->pose0; e->intr = e->intr0; e->pt = e->pt0; e->traj.clear();
    return BA_OK;
}

int ba_solve(ba_engine* e, ba_summary* out) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    const int rc = e->opt.lm_policy == BA_POLICY_G2O ? lm_solve_g2o(e) : lm_solve(e);
    if (out) *out = e->summary;
    return rc;
}

int ba_get_state(ba_engine* e, double* pose, double* intr, double* pt) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    if (pose) std::copy(e->pose.begin(), e->pose.end(), pose);
    if (intr) std::copy(e->intr.begin(), e->intr.end(), intr);
    if (pt) std::copy(e->pt.begin(), e->pt.end(), pt);
    return BA_OK;
}

int ba_get_trajectory(ba_engine* e, ba_iter_record* out, int32_t cap, int32_t* n) {
    if (!e || !n) return BA_ERR_INVALID_ARGUMENT;
    *n = int32_t(e->traj.size());
    if (out) for (int i = 0; i < std::min<int>(cap, *n); ++i) out[i] = e->traj[i];
    return BA_OK;
}

int ba_get_structure(ba_engine* e, int32_t* perm_by_cam, int32_t* cam_off, int32_t* pt_off, int32_t* schur_pairs, int64_t pairs_cap, int64_t* n_pairs) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    if (perm_by_cam) std::copy(e->perm_by_cam.begin(), e->perm_by_cam.end(), perm_by_cam);
    if (cam_off) std::copy(e->cam_off.begin(), e->cam_off.end(), cam_off);
    if (pt_off) std::copy(e->pt_off.begin(), e->pt_off.end(), pt_off);
    if (n_pairs || schur_pairs) {   // brute force: mark every unordered pair of free poses that co-observe a point
        const int Nc = e->Nc;
        vector<uint8_t> mark(size_t(Nc) * Nc, 0);
        for (int p = 0; p < e->Np; ++p)
            for (int a = e->pt_off[p]; a < e->pt_off[p + 1]; ++a)
                for (int b = e->pt_off[p]; b < e->pt_off[p + 1]; ++b) {
                    const int i = e->obs_cam[a], j = e->obs_cam[b];
                    if (i < j && e->cam_free[i] && e->cam_free[j]) mark[size_t(i) * Nc + j] = 1;
                }
        int64_t n = 0;
        for (int i = 0; i < Nc; ++i) for (int j = i + 1; j < Nc; ++j) if (mark[size_t(i) * Nc + j]) {
            if (schur_pairs && n < pairs_cap) { schur_pairs[2 * n] = i; schur_pairs[2 * n + 1] = j; }
            ++n;
        }
        if (n_pairs) *n_pairs = n;
    }
    return BA_OK;
}

int ba_evaluate(ba_engine* e, double* cost, double* r, double* Jc, double* Jp, double* Ji, double* g) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    const int No = e->No, ni = e->ni;
    vector<double> tr(size_t(2) * No), tJc(size_t(12) * No), tJp(size_t(6) * No), tJi(size_t(2) * ni * No);
    double c = 0.0;
    const bool ok = evaluate(e, e->pose.data(), e->intr.data(), e->pt.data(), true, &c, tr.data(), tJc.data(), tJp.data(), tJi.data(), false);
    if (cost) *cost = c;
    if (r) std::copy(tr.begin(), tr.end(), r);
    if (Jc) std::copy(tJc.begin(), tJc.end(), Jc);
    if (Jp) std::copy(tJp.begin(), tJp.end(), Jp);
    if (Ji) for (int o = 0; o < No; ++o) for (int q = 0; q < 2; ++q) for (int k = 0; k < 4; ++k) Ji[(2 * o + q) * 4 + k] = k < ni ? tJi[(2 * o + q) * ni + k] : 0.0;
    if (g) {
        vector<double> gc(size_t(6) * e->Nc), gi(size_t(ni) * e->Ng), gp(size_t(3) * e->Np);
        gradient(e, tr.data(), tJc.data(), tJp.data(), tJi.data(), gc.data(), gi.data(), gp.data());
        double* w = g;
        w = std::copy(gc.begin(), gc.end(), w);
        for (int gidx = 0; gidx < e->Ng; ++gidx) for (int k = 0; k < 4; ++k) *w++ = k < ni ? gi[ni * gidx + k] : 0.0;
        std::copy(gp.begin(), gp.end(), w);
    }
    return ok ? BA_OK : fail(e, BA_ERR_EVAL_FAILED, "non-finite residual or Jacobian");
}

int ba_set_profiling(ba_engine* e, int32_t) { return e ? BA_OK : BA_ERR_INVALID_ARGUMENT; }
int ba_set_profiling_stride(ba_engine* e, int32_t stride) { return (e && stride >= 1) ? BA_OK : BA_ERR_INVALID_ARGUMENT; }
int ba_get_kernel_stats(ba_engine* e, ba_kernel_stat*, int32_t, int32_t* n) { if (!e || !n) return BA_ERR_INVALID_ARGUMENT; *n = 0; return BA_OK; }
int ba_linear_solver_in_use(ba_engine* e, int32_t* solver) { if (!e || !solver) return BA_ERR_INVALID_ARGUMENT; *solver = e->opt.linear_solver == BA_LINEAR_PCG ? BA_LINEAR_PCG : BA_LINEAR_EXACT; return BA_OK; }
int ba_shard_points(int32_t, int32_t, const int32_t*, int32_t, int32_t*) { return BA_ERR_UNSUPPORTED; }   // engine-side host logic; the oracle is single-process
int ba_nccl_unique_id(void*) { return BA_ERR_UNSUPPORTED; }
int ba_get_execution_info(ba_engine* e, int64_t* info, int32_t cap, int32_t* n) {
    if (!e || !n) return BA_ERR_INVALID_ARGUMENT;
    *n = 8;
    if (info) for (int i = 0; i < std::min<int>(cap, 8); ++i) info[i] = 0;
    return BA_OK;
}
int ba_cost_noalloc(int32_t, void*, int32_t, int32_t, double, int32_t, int32_t, int32_t, int32_t, const double*, const double*, const int32_t*,
                    const double*, const int32_t*, const int32_t*, const double*, double*, int32_t*) {
    return fail(nullptr, BA_ERR_UNSUPPORTED, "ba_cost_noalloc is a device entry point of the CUDA engine; the CPU oracle has no device");
}
int ba_get_io_bytes(int64_t* h2d, int64_t* d2h) { if (h2d) *h2d = 0; if (d2h) *d2h = 0; return BA_OK; }
const char* ba_last_error(const ba_engine* e) { return e ? e->err.c_str() : g_create_error.c_str(); }
const char* ba_backend(void) { return "cpu-oracle"; }
void ba_destroy(ba_engine* e) { delete e; }

}  // extern "C"
