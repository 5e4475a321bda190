// ba_dense.cuh — direct solve of the reduced camera system (ba_linear_solver EXACT) for small camera counts.
//
// The reference configures SPARSE_SCHUR (BundleAdjuster.h:172-173): Ceres forms S = F^T F + Dc^2 - F^T E (E^T E + Dp^2)^-1 E^T F and
// factorises it.  For the problems the reference actually runs (2..~300 images, SequentialActuator.h:138-229) S has at most a few
// thousand rows, PCG needs hundreds of 4-launch iterations per LM iteration and every one of them is launch-latency-bound
// (tools/small_regime.py: 6 us per launch, 25 us per PCG iteration on a 30-camera map).  Here S is formed densely and factorised:
//
//   k_dense_wu      per observation   W_o = Jc_o^T Jp_o (6x3), U_o = W_o Hinv_p                                    (thread per obs)
//   k_dense_point_g per point         G_p = sum Ji_o^T Jp_o (4x3), G_p Hinv_p, corner partials (free intrinsics)  (thread per point)
//   k_dense_init    diag blocks Hcc + Dc^2, padding diagonal, the right-hand side as an extra row, the corner block
//   k_dense_pairs   block (ci, cj) -= sum over co-observing pairs U_i W_j^T, static by-block pair lists, no atomics (warp per block)
//   k_dense_border  intrinsics rows against every camera (free intrinsics, one shared Camera)                       (warp per camera)
//   k_dense_chol    cooperative grid: blocked right-looking Cholesky (32x32 tiles) of the matrix augmented with b^T as its last row
//                   (the forward substitution L y = b falls out of the panel solves), then L^T x = y by CTA 0.
//
// Every sum has a fixed order (pair lists are built on the host in point order, lanes take fixed strides, shuffles form a fixed
// tree), so the solve is bit-reproducible run to run.
#pragma once
#include <cooperative_groups.h>

#include "ba_kernels.cuh"

namespace bak {

constexpr int DT = 32;            // Cholesky tile
constexpr int DTP = DT + 1;       // padded shared-memory row

struct DenseView {
    int n;                 // nf = 6 Nc (+ 4 with free intrinsics)
    int T;                 // tile rows of the matrix, ceil(n / 32); tile row T is the right-hand side
    int ld;                // leading dimension of S (doubles) = 32 T
    double* S;             // [(32 (T + 1))][ld] row-major, lower triangle used; row 32 T holds b^T, then y^T
    double* WU;            // [No][36]: U (6x3 row-major), W (6x3 row-major)
    double* GH;            // [Np][24]: G (4x3), G Hinv (4x3)            (free intrinsics)
    double* corner_part;   // [corner_blocks][16]                        (free intrinsics)
    int corner_blocks;
    int nblk;
    const int* blk_off;    // [nblk + 1] into pairs
    const int2* blk_key;   // [nblk] (ci, cj), ci >= cj
    const int2* pairs;     // (observation i, observation j): block (cam i, cam j) -= U_i W_j^T
};

__global__ void __launch_bounds__(128) k_dense_wu(View V, DenseView D) {
    const int o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= V.No) return;
    double2 jc[6], jp[3];
#pragma unroll
    for (int k = 0; k < 6; ++k) jc[k] = V.J2[size_t(k) * V.ldo + o];
#pragma unroll
    for (int a = 0; a < 3; ++a) jp[a] = V.J2[size_t(6 + a) * V.ldo + o];
    const double* h = V.Hinv + 6 * size_t(V.obs_pt[o]);
    const double H[3][3] = {{h[0], h[1], h[2]}, {h[1], h[3], h[4]}, {h[2], h[4], h[5]}};
    double2* out = reinterpret_cast<double2*>(D.WU + 36 * size_t(o));
    double w[6][3], u[6][3];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
#pragma unroll
        for (int a = 0; a < 3; ++a) w[k][a] = jc[k].x * jp[a].x + jc[k].y * jp[a].y;
#pragma unroll
        for (int a = 0; a < 3; ++a) u[k][a] = w[k][0] * H[0][a] + w[k][1] * H[1][a] + w[k][2] * H[2][a];
    }
    const double* uf = &u[0][0];
    const double* wf = &w[0][0];
#pragma unroll
    for (int k = 0; k < 9; ++k) out[k] = make_double2(uf[2 * k], uf[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 9; ++k) out[9 + k] = make_double2(wf[2 * k], wf[2 * k + 1]);
}

// Free intrinsics, one shared Camera: G_p = sum_o Ji_o^T Jp_o, G_p Hinv_p, and the corner term G_p Hinv_p G_p^T reduced per CTA.
__global__ void __launch_bounds__(128) k_dense_point_g(View V, DenseView D) {
    __shared__ double sred[32];
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    double G[4][3] = {}, GHm[4][3] = {};
    if (p < V.Np) {
        for (int o = V.pt_off[p]; o < V.pt_off[p + 1]; ++o) {
            double2 jp[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) jp[a] = V.J2[size_t(6 + a) * V.ldo + o];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double2 ji = V.J2[size_t(9 + q) * V.ldo + o];
#pragma unroll
                for (int a = 0; a < 3; ++a) G[q][a] += ji.x * jp[a].x + ji.y * jp[a].y;
            }
        }
        const double* h = V.Hinv + 6 * size_t(p);
        const double H[3][3] = {{h[0], h[1], h[2]}, {h[1], h[3], h[4]}, {h[2], h[4], h[5]}};
        double* out = D.GH + 24 * size_t(p);
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                GHm[q][a] = G[q][0] * H[0][a] + G[q][1] * H[1][a] + G[q][2] * H[2][a];
                out[3 * q + a] = G[q][a];
                out[12 + 3 * q + a] = GHm[q][a];
            }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const double v = block_sum(GHm[q][0] * G[r][0] + GHm[q][1] * G[r][1] + GHm[q][2] * G[r][2], sred);
            if (threadIdx.x == 0) D.corner_part[16 * size_t(blockIdx.x) + 4 * q + r] = v;
        }
}

// S was cleared by a memset.  Diagonal blocks = Hcc + Dc^2, padding rows get a unit diagonal, b^T goes into row 32 T, and with free
// intrinsics the corner block = Hii + Di^2 - sum of the per-CTA corner partials (fixed order).
__global__ void __launch_bounds__(256) k_dense_init(int Nc, int fi, DenseView D, const double* __restrict__ Hcc, const double* __restrict__ Hii,
                                                    const double* __restrict__ Df2, const double* __restrict__ b) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int nd = 36 * Nc;
    if (i < nd) {
        const int c = i / 36, e = i - 36 * c, r = e / 6, q = e - 6 * r;
        D.S[size_t(6 * c + r) * D.ld + 6 * c + q] = Hcc[i] + (r == q ? Df2[6 * c + r] : 0.0);
    } else if (i < nd + D.ld) {
        const int j = i - nd;
        if (j >= D.n) D.S[size_t(j) * D.ld + j] = 1.0;
        else D.S[size_t(DT) * D.T * D.ld + j] = b[j];
    } else if (fi && i < nd + D.ld + 16) {
        const int e = i - nd - D.ld, q = e / 4, r = e - 4 * q;
        double acc = 0.0;
        for (int k = 0; k < D.corner_blocks; ++k) acc += D.corner_part[16 * size_t(k) + e];
        D.S[size_t(6 * Nc + q) * D.ld + 6 * Nc + r] = Hii[e] + (q == r ? Df2[6 * Nc + q] : 0.0) - acc;
    }
}

// One warp per (ci, cj) block: lanes stride over the block's pair list, 36 accumulators each, butterfly reduction, then the block
// (owned by this warp alone) is updated in place.
__global__ void __launch_bounds__(128) k_dense_pairs(DenseView D) {
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int blk = warp; blk < D.nblk; blk += nwarps) {
        const int beg = D.blk_off[blk], end = D.blk_off[blk + 1];
        double acc[36];
#pragma unroll
        for (int k = 0; k < 36; ++k) acc[k] = 0.0;
        for (int q = beg + lane; q < end; q += 32) {
            const int2 pr = D.pairs[q];
            const double2* U2 = reinterpret_cast<const double2*>(D.WU + 36 * size_t(pr.x));
            const double2* W2 = reinterpret_cast<const double2*>(D.WU + 36 * size_t(pr.y) + 18);
            double u[18], w[18];
#pragma unroll
            for (int k = 0; k < 9; ++k) { const double2 a = U2[k], c = W2[k]; u[2 * k] = a.x; u[2 * k + 1] = a.y; w[2 * k] = c.x; w[2 * k + 1] = c.y; }
#pragma unroll
            for (int r = 0; r < 6; ++r)
#pragma unroll
                for (int c = 0; c < 6; ++c) acc[6 * r + c] += u[3 * r] * w[3 * c] + u[3 * r + 1] * w[3 * c + 1] + u[3 * r + 2] * w[3 * c + 2];
        }
#pragma unroll
        for (int k = 0; k < 36; ++k)
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], d);
        const int2 key = D.blk_key[blk];
        double* Sb = D.S + size_t(6 * key.x) * D.ld + 6 * key.y;
#pragma unroll
        for (int k = 0; k < 36; ++k)
            if (lane == (k & 31)) Sb[size_t(k / 6) * D.ld + (k % 6)] -= acc[k];
    }
}

// Free intrinsics (one shared Camera): rows 6 Nc .. 6 Nc + 3 against camera c:  sum over the camera's observations of
// Ji^T Jc - (G_p Hinv_p) W_o^T.  One warp per camera, by-camera observation list, fixed order.
__global__ void __launch_bounds__(128) k_dense_border(View V, DenseView D) {
    const int lane = threadIdx.x & 31;
    const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (c >= V.Nc) return;
    double acc[24];
#pragma unroll
    for (int k = 0; k < 24; ++k) acc[k] = 0.0;
    for (int idx = V.cam_off[c] + lane; idx < V.cam_off[c + 1]; idx += 32) {
        const int o = V.perm_by_cam[idx];
        const double* gh = D.GH + 24 * size_t(V.obs_pt[o]) + 12;
        const double* w = D.WU + 36 * size_t(o) + 18;
        double2 jc[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) jc[k] = V.J2[size_t(k) * V.ldo + o];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double2 ji = V.J2[size_t(9 + q) * V.ldo + o];
            const double g0 = gh[3 * q], g1 = gh[3 * q + 1], g2 = gh[3 * q + 2];
#pragma unroll
            for (int k = 0; k < 6; ++k) acc[6 * q + k] += ji.x * jc[k].x + ji.y * jc[k].y - (g0 * w[3 * k] + g1 * w[3 * k + 1] + g2 * w[3 * k + 2]);
        }
    }
#pragma unroll
    for (int k = 0; k < 24; ++k)
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], d);
#pragma unroll
    for (int k = 0; k < 24; ++k)
        if (lane == k) D.S[size_t(6 * V.Nc + k / 6) * D.ld + 6 * c + (k % 6)] = acc[k];
}

// ---- blocked Cholesky of the augmented matrix, forward substitution folded in, then back substitution -----------------------------
namespace cg = cooperative_groups;

// In-place Cholesky of the lower triangle of a 32x32 shared-memory tile (row stride DTP) by one warp, lane = row.
__device__ __forceinline__ bool potrf32(double* L, int lane) {
    bool ok = true;
    for (int j = 0; j < DT; ++j) {
        const double d = L[j * DTP + j];
        if (!(d > 0.0)) ok = false;   // also catches NaN
        const double dj = sqrt(d);
        double l = 0.0;
        if (lane > j) { l = L[lane * DTP + j] / dj; L[lane * DTP + j] = l; }
        else if (lane == j) L[j * DTP + j] = dj;
        __syncwarp();
#pragma unroll 4
        for (int c = j + 1; c <= lane; ++c) L[lane * DTP + c] -= l * L[c * DTP + j];
        __syncwarp();
    }
    return ok;
}

// Zt[m][c] = (L^-1)[c][m] for the lower-triangular tile L, one warp, lane = column m of the inverse (zero above the diagonal).
__device__ __forceinline__ void trinv32_t(const double* L, double* Zt, int lane) {
    double* z = Zt + lane * DTP;   // z[r] = (L^-1)[r][lane]
    for (int r = 0; r < lane; ++r) z[r] = 0.0;
    z[lane] = 1.0 / L[lane * DTP + lane];
    for (int r = lane + 1; r < DT; ++r) {
        double s0 = 0.0, s1 = 0.0;
        int m = lane;
        for (; m + 1 < r; m += 2) { s0 += L[r * DTP + m] * z[m]; s1 += L[r * DTP + m + 1] * z[m + 1]; }
        if (m < r) s0 += L[r * DTP + m] * z[m];
        z[r] = -(s0 + s1) / L[r * DTP + r];
    }
}

// 256 threads move one 32x32 tile: thread t owns row t>>3, columns 4 (t&7) .. +3.
__device__ __forceinline__ void tile_load(const double* __restrict__ g, int ld, double* s, int tid) {
    const int r = tid >> 3, c = (tid & 7) * 4;
    const double2 a = *reinterpret_cast<const double2*>(g + size_t(r) * ld + c), b = *reinterpret_cast<const double2*>(g + size_t(r) * ld + c + 2);
    s[r * DTP + c] = a.x; s[r * DTP + c + 1] = a.y; s[r * DTP + c + 2] = b.x; s[r * DTP + c + 3] = b.y;
}
__device__ __forceinline__ void tile_load_t(const double* __restrict__ g, int ld, double* s, int tid) {   // s[m][r] = g[r][m]
    const int r = tid >> 3, c = (tid & 7) * 4;
    const double2 a = *reinterpret_cast<const double2*>(g + size_t(r) * ld + c), b = *reinterpret_cast<const double2*>(g + size_t(r) * ld + c + 2);
    s[c * DTP + r] = a.x; s[(c + 1) * DTP + r] = a.y; s[(c + 2) * DTP + r] = b.x; s[(c + 3) * DTP + r] = b.y;
}
// out[q] = sum_m A[r][m] Bt[m][c + q], q = 0..3
__device__ __forceinline__ void tile_mma(const double* A, const double* Bt, int tid, double out[4]) {
    const int r = tid >> 3, c = (tid & 7) * 4;
    out[0] = out[1] = out[2] = out[3] = 0.0;
#pragma unroll 8
    for (int m = 0; m < DT; ++m) {
        const double a = A[r * DTP + m];
        const double* b = Bt + m * DTP + c;
        out[0] += a * b[0]; out[1] += a * b[1]; out[2] += a * b[2]; out[3] += a * b[3];
    }
}

__global__ void __launch_bounds__(256) k_dense_chol(DenseView D, double* __restrict__ x, PcgState2* __restrict__ pcg, int* __restrict__ bad_flag) {
    __shared__ double sL[DT * DTP], sZt[DT * DTP], sA[DT * DTP], sBt[DT * DTP];
    __shared__ double sx[DT];
    __shared__ int s_ok;
    cg::grid_group grid = cg::this_grid();
    const int tid = threadIdx.x, lane = tid & 31, T = D.T, ld = D.ld;
    double* S = D.S;
    for (int k = 0; k < T; ++k) {
        // ---- panel: L_kk = chol(A_kk) (every CTA with work, redundantly), L_ik = A_ik L_kk^-T for the tile rows below, the rhs row included
        if (int(blockIdx.x) < T - k) {
            tile_load(S + size_t(DT) * k * ld + DT * k, ld, sL, tid);
            __syncthreads();
            if (tid < 32) {
                const bool ok = potrf32(sL, lane);
                trinv32_t(sL, sZt, lane);
                if (lane == 0) s_ok = ok;
            }
            __syncthreads();
            if (blockIdx.x == 0) {
                if (tid == 0 && !s_ok) atomicOr(bad_flag, 16);
                const int r = tid >> 3, c = (tid & 7) * 4;
                double* g = S + size_t(DT * k + r) * ld + DT * k + c;
#pragma unroll
                for (int q = 0; q < 4; ++q) if (c + q <= r) g[q] = sL[r * DTP + c + q];
            }
            for (int i = k + 1 + blockIdx.x; i <= T; i += gridDim.x) {
                double* g = S + size_t(DT) * i * ld + DT * k;
                __syncthreads();
                tile_load(g, ld, sA, tid);
                __syncthreads();
                double out[4];
                tile_mma(sA, sZt, tid, out);
                const int r = tid >> 3, c = (tid & 7) * 4;
                *reinterpret_cast<double2*>(g + size_t(r) * ld + c) = make_double2(out[0], out[1]);
                *reinterpret_cast<double2*>(g + size_t(r) * ld + c + 2) = make_double2(out[2], out[3]);
            }
        }
        if (k == T - 1) break;
        grid.sync();
        // ---- trailing update: A_ij -= L_ik L_jk^T for k < j <= i <= T (the never-used corner (T, T) is skipped)
        const int m = T - k, ntasks = m * (m + 1) / 2 - 1;
        for (int t = blockIdx.x; t < ntasks; t += gridDim.x) {
            int ii = int((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
            while ((ii + 1) * (ii + 2) / 2 <= t) ++ii;
            while (ii * (ii + 1) / 2 > t) --ii;
            const int jj = t - ii * (ii + 1) / 2;
            const int i = k + 1 + ii, j = k + 1 + jj;
            __syncthreads();
            tile_load(S + size_t(DT) * i * ld + DT * k, ld, sA, tid);
            tile_load_t(S + size_t(DT) * j * ld + DT * k, ld, sBt, tid);
            __syncthreads();
            double out[4];
            tile_mma(sA, sBt, tid, out);
            const int r = tid >> 3, c = (tid & 7) * 4;
            double* g = S + size_t(DT * i + r) * ld + DT * j + c;
            double2 a = *reinterpret_cast<double2*>(g), b = *reinterpret_cast<double2*>(g + 2);
            a.x -= out[0]; a.y -= out[1]; b.x -= out[2]; b.y -= out[3];
            *reinterpret_cast<double2*>(g) = a; *reinterpret_cast<double2*>(g + 2) = b;
        }
        grid.sync();
    }
    grid.sync();
    if (blockIdx.x != 0) return;
    // ---- L^T x = y, y^T = row 32 T; one CTA, block rows from the bottom, lazy update of the part of y above
    double* y = S + size_t(DT) * T * ld;
    for (int kb = T - 1; kb >= 0; --kb) {
        __syncthreads();
        tile_load(S + size_t(DT) * kb * ld + DT * kb, ld, sL, tid);
        __syncthreads();
        if (tid < 32) {
            double yv = y[DT * kb + lane];
            const double rinv = 1.0 / sL[lane * DTP + lane];
            for (int c = DT - 1; c >= 0; --c) {
                const double xc = __shfl_sync(0xffffffffu, yv * rinv, c);
                if (lane == c) yv = xc;
                else if (lane < c) yv -= sL[c * DTP + lane] * xc;
            }
            sx[lane] = yv;
            if (DT * kb + lane < D.n) x[DT * kb + lane] = yv;
        }
        __syncthreads();
        const double* Lrow = S + size_t(DT) * kb * ld;
        for (int j = tid; j < DT * kb; j += blockDim.x) {
            double acc = 0.0;
#pragma unroll 8
            for (int r = 0; r < DT; ++r) acc += Lrow[size_t(r) * ld + j] * sx[r];
            y[j] -= acc;
        }
    }
    if (tid == 0) { pcg->k = 0; pcg->done = 1; pcg->rel = 0.0; }
}

}  // namespace bak
