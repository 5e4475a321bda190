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
    // work items: a block's pair list cut into chunks of <= PAIR_CHUNK pairs, one warp per item
    int nitems;
    const int4* items;     // (block, first pair, end pair, first item of the block)
    const int* blk_item_off;   // [nblk + 1]
    double* item_part;     // [nitems][36] partial sums of multi-item blocks
    int* blk_ticket;       // [nblk] arrival counters, zero between launches (the last warp to arrive sums the partials in item order and resets)
    double* Zinv;          // [T][32][32] transposed inverses of the diagonal factor blocks: Zinv[k][c][m] = (L_kk^-1)[m][c]... see k_dense_chol
};
constexpr int PAIR_CHUNK = 64;

// Dataflow Cholesky (k_dense_chol_df): tiles in column-major order, one flag per tile
struct CholFlow {
    int ntasks;
    const int2* tasks;     // (i, j), j <= i <= T, column by column; CTA b runs tasks b, b + gridDim, ... in this order
    int* ready;            // [(T + 1)][(T + 1)]: == epoch once L_ij (or, on the diagonal, the inverse of L_jj in Zinv) is in memory
    int* abort;            // set by a wait that timed out: every later wait returns at once
    int epoch;             // solve counter, so the flags never need clearing
};

__global__ void __launch_bounds__(128) k_dense_wu(View V, DenseView D) {
    const int o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= V.No) return;
    double2 jc[6], jp[3];
    load_jc(V, V.J2 + o, jc);
    load_jp(V, V.J2 + o, jp);
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
            double2 jp[3], jiq[4];
            load_jp(V, V.J2 + o, jp);
            load_ji(V, V.J2 + o, jiq);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double2 ji = jiq[q];
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
    const int nd = 36 * Nc, cbase = (nd + D.ld + 31) / 32 * 32;
    if (i < nd) {
        const int c = i / 36, e = i - 36 * c, r = e / 6, q = e - 6 * r;
        D.S[size_t(6 * c + r) * D.ld + 6 * c + q] = Hcc[i] + (r == q ? Df2[6 * c + r] : 0.0);
    } else if (i < nd + D.ld) {
        const int j = i - nd;
        if (j >= D.n) D.S[size_t(j) * D.ld + j] = 1.0;
        else D.S[size_t(DT) * D.T * D.ld + j] = b[j];
    } else if (fi && i >= cbase && i < cbase + 16 * 32) {   // one warp per corner entry: lanes stride over the per-CTA partials, fixed tree
        const int e = (i - cbase) >> 5, lane = (i - cbase) & 31, q = e / 4, r = e - 4 * q;
        double acc = 0.0;
        for (int k = lane; k < D.corner_blocks; k += 32) acc += D.corner_part[16 * size_t(k) + e];
        acc = warp_sum(acc);
        if (lane == 0) D.S[size_t(6 * Nc + q) * D.ld + 6 * Nc + r] = Hii[e] + (q == r ? Df2[6 * Nc + q] : 0.0) - acc;
    }
}

// One warp per work item (a chunk of <= PAIR_CHUNK pairs of one (ci, cj) block): 36 accumulators per lane, butterfly reduction.  A
// block with one item is updated in place by its warp; otherwise the partial goes to item_part and the LAST warp of the block to
// arrive (ticket counter) adds the partials in item order — the sum does not depend on which warp that was.
__global__ void __launch_bounds__(128) k_dense_pairs(DenseView D) {
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int item = warp; item < D.nitems; item += nwarps) {
        const int4 it = D.items[item];
        double acc[36];
#pragma unroll
        for (int k = 0; k < 36; ++k) acc[k] = 0.0;
        for (int q = it.y + lane; q < it.z; q += 32) {
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
        const int blk = it.x;
        const int nit = D.blk_item_off[blk + 1] - D.blk_item_off[blk];
        const int2 key = D.blk_key[blk];
        double* Sb = D.S + size_t(6 * key.x) * D.ld + 6 * key.y;
        if (nit == 1) {
#pragma unroll
            for (int k = 0; k < 36; ++k)
                if (lane == (k & 31)) Sb[size_t(k / 6) * D.ld + (k % 6)] -= acc[k];
            continue;
        }
        double* mine = D.item_part + 36 * size_t(item);
#pragma unroll
        for (int k = 0; k < 36; ++k)
            if (lane == (k & 31)) __stcg(mine + k, acc[k]);
        __threadfence();
        __syncwarp();
        int last = 0;
        if (lane == 0) {
            last = atomicAdd(D.blk_ticket + blk, 1) == nit - 1;
            if (last) D.blk_ticket[blk] = 0;
        }
        last = __shfl_sync(0xffffffffu, last, 0);
        if (!last) continue;
        __threadfence();
        const double* part = D.item_part + 36 * size_t(it.w);
        double s0 = 0.0, s1 = 0.0;
        for (int i = 0; i < nit; ++i) {
            s0 += __ldcg(part + 36 * size_t(i) + lane);
            if (lane < 4) s1 += __ldcg(part + 36 * size_t(i) + 32 + lane);
        }
        Sb[size_t(lane / 6) * D.ld + (lane % 6)] -= s0;
        if (lane < 4) Sb[size_t((32 + lane) / 6) * D.ld + ((32 + lane) % 6)] -= s1;
    }
}

// Free intrinsics (one shared Camera): rows 6 Nc .. 6 Nc + 3 against camera c:  sum over the camera's observations of
// Ji^T Jc - (G_p Hinv_p) W_o^T.  One CTA per camera over its by-camera observation list, fixed order.
__global__ void __launch_bounds__(128) k_dense_border(View V, DenseView D) {
    __shared__ double sm[4][24];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int c = blockIdx.x;
    double acc[24];
#pragma unroll
    for (int k = 0; k < 24; ++k) acc[k] = 0.0;
    for (int idx = V.cam_off[c] + threadIdx.x; idx < V.cam_off[c + 1]; idx += blockDim.x) {
        const int o = V.perm_by_cam[idx];
        const double* gh = D.GH + 24 * size_t(V.obs_pt[o]) + 12;
        const double* w = D.WU + 36 * size_t(o) + 18;
        double2 jc[6], jiq[4];
        load_jc(V, V.J2 + o, jc);
        load_ji(V, V.J2 + o, jiq);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double2 ji = jiq[q];
            const double g0 = gh[3 * q], g1 = gh[3 * q + 1], g2 = gh[3 * q + 2];
#pragma unroll
            for (int k = 0; k < 6; ++k) acc[6 * q + k] += ji.x * jc[k].x + ji.y * jc[k].y - (g0 * w[3 * k] + g1 * w[3 * k + 1] + g2 * w[3 * k + 2]);
        }
    }
#pragma unroll
    for (int k = 0; k < 24; ++k) { const double v = warp_sum(acc[k]); if (lane == 0) sm[wid][k] = v; }
    __syncthreads();
    if (threadIdx.x < 24) {
        const int k = threadIdx.x;
        D.S[size_t(6 * V.Nc + k / 6) * D.ld + 6 * c + (k % 6)] = (sm[0][k] + sm[1][k]) + (sm[2][k] + sm[3][k]);
    }
}

// ---- blocked Cholesky of the augmented matrix, forward substitution folded in, then back substitution -----------------------------
namespace cg = cooperative_groups;

// Cholesky of the lower triangle of the 32x32 shared-memory tile L (row stride DTP) AND the transposed inverse of the factor, by one
// warp with its row in registers (lane = row; every loop is unrolled, so a[] never leaves the register file):
//   column j:  r = rsqrt(a_jj) from lane j, the column is scaled by r (so L_jj = a_jj r), published through `col` (one STS, broadcast
//   LDS.128 reads), and every lane folds it into the rest of its row.  One rsqrt, no division, no shared-memory dependency chain.
// Then lane c solves L z = e_c for column c of L^-1 with z in registers and L read back through broadcast loads (Lb, row stride DT + 2
// so that rows are 16-byte aligned); 1 / L_jj is the rsqrt already computed.  Writes L back to `L` and Zt[m][c] = (L^-1)[c][m].
constexpr int DTB = DT + 2;
__device__ __forceinline__ bool potrf_trinv32(double* L, double* Zt, double* Lb, double* col, double* rinv, int lane) {
    double a[DT];
#pragma unroll
    for (int c = 0; c < DT; ++c) a[c] = L[lane * DTP + c];
    bool ok = true;
#pragma unroll
    for (int j = 0; j < DT; ++j) {
        const double d = __shfl_sync(0xffffffffu, a[j], j);
        if (!(d > 0.0)) ok = false;   // also catches NaN
        const double r = rsqrt(d);
        a[j] *= r;                    // lane j: L_jj = d r; lanes below: L_rj
        if (lane == j) rinv[j] = r;
        if (j + 1 < DT) {
            col[lane] = a[j];
            __syncwarp();
#pragma unroll
            for (int c = j + 1; c < DT; ++c) a[c] -= a[j] * col[c];   // lanes >= c hold real entries; the rest is never read
            __syncwarp();
        }
    }
#pragma unroll
    for (int c = 0; c < DT; ++c) {
        const double v = c <= lane ? a[c] : 0.0;
        L[lane * DTP + c] = v;
        Lb[lane * DTB + c] = v;
    }
    __syncwarp();
    double z[DT];   // z[r] = (L^-1)[r][lane]
#pragma unroll
    for (int r = 0; r < DT; ++r) {
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
        for (int m = 0; m < r; ++m) {
            const double t = Lb[r * DTB + m] * z[m];
            if ((m & 3) == 0) s0 += t; else if ((m & 3) == 1) s1 += t; else if ((m & 3) == 2) s2 += t; else s3 += t;
        }
        const double ri = rinv[r];
        z[r] = r < lane ? 0.0 : (r == lane ? ri : -((s0 + s1) + (s2 + s3)) * ri);
    }
#pragma unroll
    for (int r = 0; r < DT; ++r) Zt[lane * DTP + r] = z[r];
    return ok;
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
    __shared__ __align__(16) double sLb[DT * DTB];
    __shared__ __align__(16) double scol[DT];
    __shared__ double sx[DT], srinv[DT];
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
                const bool ok = potrf_trinv32(sL, sZt, sLb, scol, srinv, lane);
                if (lane == 0) s_ok = ok;
            }
            __syncthreads();
            if (blockIdx.x == 0) {   // keep L_kk^-1 (transposed) for the back substitution; L_kk itself is not needed again
                if (tid == 0 && !s_ok) atomicOr(bad_flag, 16);
                const int r = tid >> 3, c = (tid & 7) * 4;
                double* g = D.Zinv + size_t(DT * DT) * k + DT * r + c;
#pragma unroll
                for (int q = 0; q < 4; ++q) g[q] = sZt[r * DTP + c + q];
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
    // ---- L^T x = y, y^T = row 32 T; one CTA, block rows from the bottom: x_k = L_kk^-T y_k with the stored inverse (no sequential
    // substitution), then the part of y above is updated with block row k of L
    double* y = S + size_t(DT) * T * ld;
    for (int kb = T - 1; kb >= 0; --kb) {
        __syncthreads();
        tile_load(D.Zinv + size_t(DT * DT) * kb, DT, sZt, tid);   // sZt[c][m] = (L_kk^-1)[m][c]
        if (tid < DT) sx[tid] = y[DT * kb + tid];
        __syncthreads();
        if (tid < DT) {
            double a0 = 0.0, a1 = 0.0;
#pragma unroll
            for (int m = 0; m < DT; m += 2) { a0 += sZt[tid * DTP + m] * sx[m]; a1 += sZt[tid * DTP + m + 1] * sx[m + 1]; }
            const double xv = a0 + a1;
            __syncwarp();
            sx[tid] = xv;
            if (DT * kb + tid < D.n) x[DT * kb + tid] = xv;
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

// ---- the same factorisation as a dataflow over tiles: no grid-wide barrier inside the factorisation -------------------------------
// Task (i, j) is LEFT-looking: acc = sum_{k<j} L_ik L_jk^T is built in registers as the operand tiles become ready (one flag per
// tile, st.release.gpu / ld.acquire.gpu), then  diagonal: factorise + invert A_jj - acc, publish the inverse;  below the diagonal:
// L_ij = (A_ij - acc) L_jj^-T, publish.  Tasks are dealt to the CTAs round-robin in column-major order and every CTA works through
// its list in order; that order is a topological order of the dependency graph and all CTAs are co-resident (cooperative launch), so
// the earliest unfinished task can always run: no deadlock.  The critical path per column is factorise+invert, one panel tile, one
// rank-32 update (vs two grid barriers plus the same work before).  A wait that exceeds ~2 s sets `abort` and the solve fails cleanly.
__device__ __forceinline__ int ld_acquire_gpu(const int* p) { int v; asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_release_gpu(int* p, int v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

__device__ __forceinline__ bool flow_wait(const int* flag, int epoch, int* abort) {
    if (ld_acquire_gpu(flag) == epoch) return true;
    const long long t0 = clock64();
    for (;;) {
        if (ld_acquire_gpu(flag) == epoch) return true;
        if (*reinterpret_cast<volatile int*>(abort)) return false;
        if (clock64() - t0 > 4000000000ll) { atomicExch(abort, 1); return false; }
    }
}

__device__ __forceinline__ void tile_load_cg(const double* __restrict__ g, int ld, double* s, int tid) {   // L2 loads: operands written by other SMs
    const int r = tid >> 3, c = (tid & 7) * 4;
    const double2 a = __ldcg(reinterpret_cast<const double2*>(g + size_t(r) * ld + c)), b = __ldcg(reinterpret_cast<const double2*>(g + size_t(r) * ld + c + 2));
    s[r * DTP + c] = a.x; s[r * DTP + c + 1] = a.y; s[r * DTP + c + 2] = b.x; s[r * DTP + c + 3] = b.y;
}
__device__ __forceinline__ void tile_load_t_cg(const double* __restrict__ g, int ld, double* s, int tid) {   // s[m][r] = g[r][m]
    const int r = tid >> 3, c = (tid & 7) * 4;
    const double2 a = __ldcg(reinterpret_cast<const double2*>(g + size_t(r) * ld + c)), b = __ldcg(reinterpret_cast<const double2*>(g + size_t(r) * ld + c + 2));
    s[c * DTP + r] = a.x; s[(c + 1) * DTP + r] = a.y; s[(c + 2) * DTP + r] = b.x; s[(c + 3) * DTP + r] = b.y;
}

// Finalised tiles are stored TRANSPOSED in global memory (tile (i, j) holds L_ij^T, element (m, r) at [m * ld + r]) so that both
// operands of L_ik L_jk^T arrive in the [m][row] layout the product wants with straight 16-byte cp.async copies (L2 -> shared, no L1).
// Product: warp w takes m in [8 (w>>1), +8) and columns [16 (w&1), +16); lane (tr, tc) = (lane>>2, lane&3) owns rows 4tr..+3 and
// columns +4tc..+3: 4 LDS.128 per 16 DFMA (the first version, one row x four columns per thread over all m, issued one LDS per DFMA
// and was shared-memory-bandwidth-bound: 2.3 us per 32x32x32 product, tools/probe/chol_probe.cu).  The four m-groups keep their
// partial sums in registers over the whole task and are added through shared memory once, at the end.
constexpr int DTQ = DT + 2;             // operand row stride: rows 16-byte aligned
constexpr int TQ = DT * DTQ;

__device__ __forceinline__ void tile_cp_async(const double* __restrict__ g, int ld, double* s, int tid) {   // s[m][r] <- g[m * ld + r], 256 threads
    const int m = tid >> 3, r4 = (tid & 7) * 4;
    cp_async16(s + m * DTQ + r4, g + size_t(m) * ld + r4);
    cp_async16(s + m * DTQ + r4 + 2, g + size_t(m) * ld + r4 + 2);
}
__device__ __forceinline__ void cp_async_wait_1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// acc[4][4] += sum over this warp's 8 m of At[m][rows] * Bt[m][cols]
__device__ __forceinline__ void tile_fma(const double* At, const double* Bt, int warp, int lane, double acc[16]) {
    const int m0 = 8 * (warp >> 1), r0 = 4 * (lane >> 2), c0 = 16 * (warp & 1) + 4 * (lane & 3);
#pragma unroll
    for (int mm = 0; mm < 8; ++mm) {
        const double2 a01 = *reinterpret_cast<const double2*>(At + (m0 + mm) * DTQ + r0), a23 = *reinterpret_cast<const double2*>(At + (m0 + mm) * DTQ + r0 + 2);
        const double2 b01 = *reinterpret_cast<const double2*>(Bt + (m0 + mm) * DTQ + c0), b23 = *reinterpret_cast<const double2*>(Bt + (m0 + mm) * DTQ + c0 + 2);
        const double a[4] = {a01.x, a01.y, a23.x, a23.y}, b[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[4 * p + q] += a[p] * b[q];
    }
}
// partial sums of the 4 m-groups -> red[g][row][col] (row stride DTQ)
__device__ __forceinline__ void tile_spill(double* red, int warp, int lane, const double acc[16]) {
    const int r0 = 4 * (lane >> 2), c0 = 16 * (warp & 1) + 4 * (lane & 3);
    double* base = red + (warp >> 1) * TQ;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        *reinterpret_cast<double2*>(base + (r0 + p) * DTQ + c0) = make_double2(acc[4 * p], acc[4 * p + 1]);
        *reinterpret_cast<double2*>(base + (r0 + p) * DTQ + c0 + 2) = make_double2(acc[4 * p + 2], acc[4 * p + 3]);
    }
}

__global__ void __launch_bounds__(256) k_dense_chol_df(DenseView D, CholFlow F, double* __restrict__ x, PcgState2* __restrict__ pcg, int* __restrict__ bad_flag) {
    // shared memory (46.6 KB static): 4 operand buffers [32][34] (A0 B0 A1 B1; as one block also the 4 x [32][34] reduction scratch,
    // and on the diagonal path the tile being factorised, its broadcast copy and the transposed inverse), plus Zt [32][34]
    __shared__ __align__(16) double sm[5 * TQ];
    __shared__ __align__(16) double scol[DT];
    __shared__ double sx[DT], srinv[DT];
    __shared__ int s_go;
    double* const sZ = sm + 4 * TQ;
    cg::grid_group grid = cg::this_grid();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, T = D.T, ld = D.ld, W = T + 1;
    const int r = tid >> 3, c = (tid & 7) * 4;
    double* S = D.S;
    bool alive = true;
    for (int t = blockIdx.x; t < F.ntasks && alive; t += gridDim.x) {
        const int2 ij = F.tasks[t];
        const int i = ij.x, j = ij.y;
        const bool diag = i == j;
        double acc[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) acc[q] = 0.0;
        // ---- acc = sum_{k<j} L_ik L_jk^T: operands double-buffered with cp.async, the copy of step k+1 overlaps the product of step k
        if (j > 0) {
            if (tid == 0) s_go = flow_wait(F.ready + i * W, F.epoch, F.abort) && (diag || flow_wait(F.ready + j * W, F.epoch, F.abort));
            __syncthreads();
            if (!s_go) { alive = false; break; }
            tile_cp_async(S + size_t(DT) * i * ld, ld, sm, tid);
            if (!diag) tile_cp_async(S + size_t(DT) * j * ld, ld, sm + TQ, tid);
            cp_async_commit();
        }
        for (int k = 0; k < j; ++k) {
            const bool more = k + 1 < j;
            if (more && tid == 0) s_go = flow_wait(F.ready + i * W + k + 1, F.epoch, F.abort) && (diag || flow_wait(F.ready + j * W + k + 1, F.epoch, F.abort));
            __syncthreads();                       // flags of step k+1 seen; everyone is done with the buffers of step k-1
            if (more && !s_go) { alive = false; break; }
            if (more) {
                double* nb = sm + 2 * ((k + 1) & 1) * TQ;
                tile_cp_async(S + size_t(DT) * i * ld + DT * (k + 1), ld, nb, tid);
                if (!diag) tile_cp_async(S + size_t(DT) * j * ld + DT * (k + 1), ld, nb + TQ, tid);
                cp_async_commit();
                cp_async_wait_1();
            } else {
                cp_async_wait_all();
            }
            __syncthreads();                       // step k's tiles are complete
            const double* At = sm + 2 * (k & 1) * TQ;
            tile_fma(At, diag ? At : At + TQ, warp, lane, acc);
        }
        if (!alive) { cp_async_wait_all(); break; }
        // ---- v = own tile (still row-major, as the formation kernels wrote it) - acc
        __syncthreads();
        tile_spill(sm, warp, lane, acc);
        double* g = S + size_t(DT * i + r) * ld + DT * j + c;
        const double2 a0 = *reinterpret_cast<const double2*>(g), a1 = *reinterpret_cast<const double2*>(g + 2);
        __syncthreads();
        double v[4];
        {
            const double* p0 = sm + r * DTQ + c;
            const double2 s00 = *reinterpret_cast<const double2*>(p0), s01 = *reinterpret_cast<const double2*>(p0 + 2);
            const double2 s10 = *reinterpret_cast<const double2*>(p0 + TQ), s11 = *reinterpret_cast<const double2*>(p0 + TQ + 2);
            const double2 s20 = *reinterpret_cast<const double2*>(p0 + 2 * TQ), s21 = *reinterpret_cast<const double2*>(p0 + 2 * TQ + 2);
            const double2 s30 = *reinterpret_cast<const double2*>(p0 + 3 * TQ), s31 = *reinterpret_cast<const double2*>(p0 + 3 * TQ + 2);
            v[0] = a0.x - ((s00.x + s10.x) + (s20.x + s30.x)); v[1] = a0.y - ((s00.y + s10.y) + (s20.y + s30.y));
            v[2] = a1.x - ((s01.x + s11.x) + (s21.x + s31.x)); v[3] = a1.y - ((s01.y + s11.y) + (s21.y + s31.y));
        }
        __syncthreads();
        if (diag) {
            double* sL = sm;                 // [32][33]
            double* sLb = sm + TQ;           // [32][34]
            double* sZt = sm + 2 * TQ;       // [32][33]: sZt[m][c] = (L^-1)[c][m]
            sL[r * DTP + c] = v[0]; sL[r * DTP + c + 1] = v[1]; sL[r * DTP + c + 2] = v[2]; sL[r * DTP + c + 3] = v[3];
            __syncthreads();
            if (tid < 32) {
                const bool ok = potrf_trinv32(sL, sZt, sLb, scol, srinv, lane);
                if (lane == 0 && !ok) atomicOr(bad_flag, 16);
            }
            __syncthreads();
            double* z = D.Zinv + size_t(DT * DT) * j + DT * r + c;
            *reinterpret_cast<double2*>(z) = make_double2(sZt[r * DTP + c], sZt[r * DTP + c + 1]);
            *reinterpret_cast<double2*>(z + 2) = make_double2(sZt[r * DTP + c + 2], sZt[r * DTP + c + 3]);
        } else {
            // L_ij = v L_jj^-T: A operand = v^T ([m][row]), B operand = Zt ([m][c]) straight from Zinv
            double* At = sm;
            At[c * DTQ + r] = v[0]; At[(c + 1) * DTQ + r] = v[1]; At[(c + 2) * DTQ + r] = v[2]; At[(c + 3) * DTQ + r] = v[3];
            if (tid == 0) s_go = flow_wait(F.ready + j * W + j, F.epoch, F.abort);
            __syncthreads();
            if (!s_go) { alive = false; break; }
            tile_cp_async(D.Zinv + size_t(DT * DT) * j, DT, sZ, tid);
            cp_async_commit();
            cp_async_wait_all();
            __syncthreads();
#pragma unroll
            for (int q = 0; q < 16; ++q) acc[q] = 0.0;
            tile_fma(At, sZ, warp, lane, acc);
            __syncthreads();
            tile_spill(sm, warp, lane, acc);
            __syncthreads();
            // transposed write-back: thread (m = r, rows c..c+3) sums the four partials of L_ij[c..c+3][m] and stores them contiguously
            double o[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double* p0 = sm + (c + q) * DTQ + r;
                o[q] = (p0[0] + p0[TQ]) + (p0[2 * TQ] + p0[3 * TQ]);
            }
            double* gt = S + size_t(DT * i + r) * ld + DT * j + c;      // element (m = r, row c + q) of the transposed tile
            *reinterpret_cast<double2*>(gt) = make_double2(o[0], o[1]);
            *reinterpret_cast<double2*>(gt + 2) = make_double2(o[2], o[3]);
        }
        __syncthreads();
        if (tid == 0) { __threadfence(); st_release_gpu(F.ready + i * W + j, F.epoch); }   // barrier, then fence + flag by one thread (cumulative)
    }
    grid.sync();
    if (blockIdx.x != 0) return;
    if (*reinterpret_cast<volatile int*>(F.abort)) { if (tid == 0) atomicOr(bad_flag, 16); return; }
    // ---- L^T x = y by CTA 0.  y_j[c] is row 0 of the (transposed) right-hand-side tile (T, j): S[(32 T + c) ld + 32 j].
    for (int kb = T - 1; kb >= 0; --kb) {
        __syncthreads();
        tile_cp_async(D.Zinv + size_t(DT * DT) * kb, DT, sZ, tid);   // sZ[c][m] = (L_kk^-1)[m][c]
        cp_async_commit();
        if (tid < DT) sx[tid] = __ldcg(S + size_t(DT * T + tid) * ld + DT * kb);
        cp_async_wait_all();
        __syncthreads();
        if (tid < DT) {
            double a0 = 0.0, a1 = 0.0;
#pragma unroll
            for (int m = 0; m < DT; m += 2) { a0 += sZ[tid * DTQ + m] * sx[m]; a1 += sZ[tid * DTQ + m + 1] * sx[m + 1]; }
            const double xv = a0 + a1;
            __syncwarp();
            sx[tid] = xv;
            if (DT * kb + tid < D.n) x[DT * kb + tid] = xv;
        }
        __syncthreads();
        // y_jt[cc] -= sum_rr L_(kb, jt)[rr][cc] x[rr]; the tile is transposed: element (rr, cc) at [(32 kb + cc) ld + 32 jt + rr]
        for (int jj = tid; jj < DT * kb; jj += blockDim.x) {
            const int jt = jj >> 5, cc = jj & 31;
            const double2* row = reinterpret_cast<const double2*>(S + size_t(DT * kb + cc) * ld + DT * jt);
            double a = 0.0, b = 0.0;
#pragma unroll
            for (int q = 0; q < DT / 2; ++q) { const double2 l = __ldcg(row + q); a += l.x * sx[2 * q]; b += l.y * sx[2 * q + 1]; }
            double* yp = S + size_t(DT * T + cc) * ld + DT * jt;
            *yp = __ldcg(yp) - (a + b);
        }
    }
    if (tid == 0) { pcg->k = 0; pcg->done = 1; pcg->rel = 0.0; }
}

}  // namespace bak
