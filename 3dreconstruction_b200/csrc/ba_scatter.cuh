// ba_scatter.cuh — atomic-free, bit-reproducible camera-side scatter for K4 (reduced RHS) and K5 (implicit Schur matvec),
// plus the multi-CTA PCG vector kernels.  
//
// Why: the point-sorted sweep produces, per observation, a 6-vector that must be added into its camera's slot.  Doing that
// with red.global.add.f64 costs 6 L2 atomics per observation (30 M per matvec on Venice) and measured 1.74 ms per matvec,
// 7 % of the HBM roofline, while the same sweep without the scatter (K7) runs at 75 %.  Here instead:
//   * a persistent grid of G CTAs; CTA g owns a fixed contiguous range of tiles and a PRIVATE accumulator acc[g][nf]
//     in global memory (L1/L2 resident), so no other CTA ever touches it;
//   * inside a tile the observations' 6-vectors are written to shared memory at their rank in the tile's camera-sorted
//     order (a static schedule computed once on the host: `slot`), the thread at each segment head sums its segment and
//     does one plain read-modify-write of acc[g][6 cam .. 6 cam + 5] — cameras are unique per tile step, steps are
//     separated by __syncthreads, so no atomics are needed;
//   * k_reduce_scratch adds the G private copies in fixed order.  Summation order is fixed => bit-reproducible.
#pragma once
#include "ba_kernels.cuh"

namespace bak {

constexpr int SCAT_CTAS_PER_SM = 2;

struct ScatterSched {
    const unsigned char* slot;   // [No] rank of the observation in its tile's camera-sorted order
    const int* spack;            // [No] per sorted position: camera | (segment length << 22) at segment heads, -1 elsewhere
    double* scratch;             // [G][ldf] private accumulators (zero on entry)
    int ldf;                     // leading dimension (>= nf, multiple of 2)
    int G;
};

template <int MODE, bool FI>
__global__ void __launch_bounds__(TILE_OBS, SCAT_CTAS_PER_SM)
k_schur_scatter(View V, ScatterSched Sch, const double* __restrict__ xin, const int* __restrict__ done_flag) {
    __shared__ double sW[TILE_OBS][3];
    __shared__ double sV[TILE_OBS][3];
    __shared__ __align__(16) double sC[TILE_OBS][6];
    __shared__ double sred[32];
    if (done_flag && *done_flag) return;
    const int G = gridDim.x, b = blockIdx.x;
    const int t0 = int((long long)V.nTiles * b / G), t1 = int((long long)V.nTiles * (b + 1) / G);
    double* __restrict__ acc = Sch.scratch + size_t(b) * Sch.ldf;
    const size_t ld = V.ldo;
    const int ioff = 6 * V.Nc;
    double yi[4] = {0, 0, 0, 0};

    for (int t = t0; t < t1; ++t) {
        const int o0 = V.tile_off[t], o1 = V.tile_off[t + 1];
        if (o1 - o0 <= TILE_OBS) {
            const int o = o0 + threadIdx.x;
            const bool act = o < o1;
            double2 jc[6], jp[3], ji[4];
            double2 u = make_double2(0.0, 0.0);
            int c = 0, p = 0, g = 0;
            if (act) {
                c = V.obs_cam[o]; p = V.obs_pt[o];
                const double2* J = V.J2 + o;
#pragma unroll
                for (int k = 0; k < 6; ++k) jc[k] = ldg2(J + k * ld);
#pragma unroll
                for (int k = 0; k < 3; ++k) jp[k] = ldg2(J + (6 + k) * ld);
                if (FI) {
                    g = V.cam_group[c];
#pragma unroll
                    for (int k = 0; k < 4; ++k) ji[k] = ldg2(J + (9 + k) * ld);
                }
                if (MODE == SWEEP_MATVEC) {
                    const double* xc = xin + 6 * c;
#pragma unroll
                    for (int k = 0; k < 6; ++k) { const double xv = __ldg(xc + k); u.x += jc[k].x * xv; u.y += jc[k].y * xv; }
                    if (FI) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) { const double xv = __ldg(xin + ioff + 4 * g + k); u.x += ji[k].x * xv; u.y += ji[k].y * xv; }
                    }
#pragma unroll
                    for (int a = 0; a < 3; ++a) sW[threadIdx.x][a] = jp[a].x * u.x + jp[a].y * u.y;
                }
            }
            double2 tt = make_double2(0.0, 0.0);
            if (MODE == SWEEP_MATVEC) {
                __syncthreads();
                const int pfirst = V.obs_pt[o0], plast = V.obs_pt[o1 - 1];
                const int pj = pfirst + threadIdx.x;
                if (pj <= plast) {
                    double W[3] = {0, 0, 0};
                    for (int i = V.pt_off[pj] - o0, e = V.pt_off[pj + 1] - o0; i < e; ++i) { W[0] += sW[i][0]; W[1] += sW[i][1]; W[2] += sW[i][2]; }
                    const double* H = V.Hinv + 6 * size_t(pj);
                    sV[threadIdx.x][0] = H[0] * W[0] + H[1] * W[1] + H[2] * W[2];
                    sV[threadIdx.x][1] = H[1] * W[0] + H[3] * W[1] + H[4] * W[2];
                    sV[threadIdx.x][2] = H[2] * W[0] + H[4] * W[1] + H[5] * W[2];
                }
                __syncthreads();
                if (act) {
                    const double* v = sV[p - pfirst];
                    tt.x = u.x - (jp[0].x * v[0] + jp[1].x * v[1] + jp[2].x * v[2]);
                    tt.y = u.y - (jp[0].y * v[0] + jp[1].y * v[1] + jp[2].y * v[2]);
                }
            } else if (act) {   // SWEEP_RHS: t = r - E vp
                const double* v = V.vp + 3 * size_t(p);
                const double2 r = V.r2[o];
                tt.x = r.x - (jp[0].x * v[0] + jp[1].x * v[1] + jp[2].x * v[2]);
                tt.y = r.y - (jp[0].y * v[0] + jp[1].y * v[1] + jp[2].y * v[2]);
            }
            // camera-side scatter: to shared memory in camera-sorted order, one owner per camera does the RMW
            if (act) {
                double* sc = sC[Sch.slot[o]];
#pragma unroll
                for (int k = 0; k < 6; ++k) sc[k] = jc[k].x * tt.x + jc[k].y * tt.y;
                if (FI) {
                    if (V.Ng > 1) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) atomicAdd(acc + ioff + 4 * g + k, ji[k].x * tt.x + ji[k].y * tt.y);
                    } else {
#pragma unroll
                        for (int k = 0; k < 4; ++k) yi[k] += ji[k].x * tt.x + ji[k].y * tt.y;
                    }
                }
            }
            __syncthreads();
            if (act) {
                const int pk = Sch.spack[o];
                if (pk >= 0) {
                    const int cam = pk & 0x3fffff, len = pk >> 22;
                    double s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0, s5 = 0;
                    for (int i = 0; i < len; ++i) {
                        const double2* row = reinterpret_cast<const double2*>(sC[threadIdx.x + i]);
                        const double2 a = row[0], bq = row[1], cq = row[2];
                        s0 += a.x; s1 += a.y; s2 += bq.x; s3 += bq.y; s4 += cq.x; s5 += cq.y;
                    }
                    // .cg (L2) accesses: the rare long-point path below updates the same private copy with L2 atomics
                    double2* dst = reinterpret_cast<double2*>(acc + 6 * size_t(cam));
                    double2 a = __ldcg(dst), bq = __ldcg(dst + 1), cq = __ldcg(dst + 2);
                    a.x += s0; a.y += s1; bq.x += s2; bq.y += s3; cq.x += s4; cq.y += s5;
                    __stcg(dst, a); __stcg(dst + 1, bq); __stcg(dst + 2, cq);
                }
            }
            __syncthreads();
        } else {
            // ---- one point with more than TILE_OBS observations (rare): two passes, atomics into the private copy ----
            const int p = V.obs_pt[o0];
            double v[3];
            if (MODE == SWEEP_MATVEC) {
                double W[3] = {0, 0, 0};
                for (int o = o0 + threadIdx.x; o < o1; o += TILE_OBS) {
                    const int c = V.obs_cam[o];
                    const double2* J = V.J2 + o;
                    double2 u = make_double2(0.0, 0.0);
#pragma unroll
                    for (int k = 0; k < 6; ++k) { const double2 j = ldg2(J + k * ld); const double xv = __ldg(xin + 6 * c + k); u.x += j.x * xv; u.y += j.y * xv; }
                    if (FI) {
                        const int g = V.cam_group[c];
#pragma unroll
                        for (int k = 0; k < 4; ++k) { const double2 j = ldg2(J + (9 + k) * ld); const double xv = __ldg(xin + ioff + 4 * g + k); u.x += j.x * xv; u.y += j.y * xv; }
                    }
#pragma unroll
                    for (int a = 0; a < 3; ++a) { const double2 j = ldg2(J + (6 + a) * ld); W[a] += j.x * u.x + j.y * u.y; }
                }
#pragma unroll
                for (int a = 0; a < 3; ++a) { const double s = block_sum(W[a], sred); if (threadIdx.x == 0) sV[0][a] = s; }
                __syncthreads();
                const double* H = V.Hinv + 6 * size_t(p);
                const double w0 = sV[0][0], w1 = sV[0][1], w2 = sV[0][2];
                v[0] = H[0] * w0 + H[1] * w1 + H[2] * w2; v[1] = H[1] * w0 + H[3] * w1 + H[4] * w2; v[2] = H[2] * w0 + H[4] * w1 + H[5] * w2;
            } else { v[0] = V.vp[3 * size_t(p)]; v[1] = V.vp[3 * size_t(p) + 1]; v[2] = V.vp[3 * size_t(p) + 2]; }
            for (int o = o0 + threadIdx.x; o < o1; o += TILE_OBS) {
                const int c = V.obs_cam[o];
                const int g = FI ? V.cam_group[c] : 0;
                const double2* J = V.J2 + o;
                double2 jc[6], ji[4];
                double2 u = make_double2(0.0, 0.0);
#pragma unroll
                for (int k = 0; k < 6; ++k) jc[k] = ldg2(J + k * ld);
                if (FI) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) ji[k] = ldg2(J + (9 + k) * ld);
                }
                if (MODE == SWEEP_MATVEC) {
#pragma unroll
                    for (int k = 0; k < 6; ++k) { const double xv = __ldg(xin + 6 * c + k); u.x += jc[k].x * xv; u.y += jc[k].y * xv; }
                    if (FI) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) { const double xv = __ldg(xin + ioff + 4 * g + k); u.x += ji[k].x * xv; u.y += ji[k].y * xv; }
                    }
                }
                double e0 = 0.0, e1 = 0.0;
#pragma unroll
                for (int a = 0; a < 3; ++a) { const double2 j = ldg2(J + (6 + a) * ld); e0 += j.x * v[a]; e1 += j.y * v[a]; }
                double2 tt;
                if (MODE == SWEEP_MATVEC) { tt.x = u.x - e0; tt.y = u.y - e1; }
                else { const double2 r = V.r2[o]; tt.x = r.x - e0; tt.y = r.y - e1; }
#pragma unroll
                for (int k = 0; k < 6; ++k) atomicAdd(acc + 6 * size_t(c) + k, jc[k].x * tt.x + jc[k].y * tt.y);
                if (FI) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const double contrib = ji[k].x * tt.x + ji[k].y * tt.y;
                        if (V.Ng > 1) atomicAdd(acc + ioff + 4 * g + k, contrib); else yi[k] += contrib;
                    }
                }
            }
            __syncthreads();
        }
    }
    if (FI && V.Ng == 1) {   // one shared intrinsics group: reduce the per-thread sums over the CTA, one plain add into the private copy
#pragma unroll
        for (int k = 0; k < 4; ++k) { const double s = block_sum(yi[k], sred); if (threadIdx.x == 0) __stcg(acc + ioff + k, __ldcg(acc + ioff + k) + s); }
    }
}


// ---------------------------------------------------------------------------------------------------------
// Multi-CTA PCG vector kernels.  One PCG iteration = k_schur_scatter<MATVEC> -> k_pcg_q -> k_pcg_xr -> k_pcg_p.
// Global dot products use per-block partials summed IN INDEX ORDER by whichever block finishes last
// (threadfence + ticket), so results do not depend on block scheduling.
constexpr int PCG_BLOCK = 128;

// y = sum_g scratch[g] (and scratch <- 0); with do_dot: q = y + Df2 p stored in y, pq = p.q, alpha = rho/pq.
// do_reduce = 0 skips the copy reduction (multi-GPU: y was already reduced and all-reduced).
__global__ void __launch_bounds__(PCG_BLOCK)
k_pcg_q(int nf, int G, int ldf, double* __restrict__ scratch, const double* __restrict__ Df2, const double* __restrict__ p,
        double* __restrict__ y, double* __restrict__ partial, PcgState2* __restrict__ st, int do_reduce, int do_dot) {
    __shared__ double sred[32];
    __shared__ bool last;
    if (do_dot && st->done) return;
    double acc = 0.0;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < nf; j += gridDim.x * blockDim.x) {
        double s = 0.0;
        if (do_reduce) {
            for (int g = 0; g < G; ++g) { double* a = scratch + size_t(g) * ldf + j; s += __ldcg(a); __stcg(a, 0.0); }
        } else s = y[j];
        if (do_dot) { s += Df2[j] * p[j]; acc += p[j] * s; }
        y[j] = s;
    }
    if (!do_dot) return;
    acc = block_sum(acc, sred);
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = acc;
        __threadfence();
        const unsigned int tk = atomicAdd(&st->ticket_q, 1u);
        last = (tk == gridDim.x - 1);
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence();
        double pq = 0.0;
        for (int i = 0; i < int(gridDim.x); ++i) pq += __ldcg(partial + i);
        if (!(pq > 0.0) || !isfinite(pq)) { st->done = 3; st->alpha = 0.0; }
        else st->alpha = st->rho / pq;
        st->ticket_q = 0;
    }
}

// One thread per pose block (and per intrinsics group): x += alpha p, r -= alpha q, z = Minv r; partials of r.z and x.(b+r);
// the last block closes the iteration: rho, beta, relative residual, stop rules.
__global__ void __launch_bounds__(PCG_BLOCK)
k_pcg_xr(int Nc, int Ng, int fi, const double* __restrict__ Minv, const double* __restrict__ b, const double* __restrict__ q,
         const double* __restrict__ p, double* __restrict__ x, double* __restrict__ r, double* __restrict__ z,
         double* __restrict__ partial, PcgState2* __restrict__ st, PcgParams pp) {
    __shared__ double sred[32];
    __shared__ bool last;
    if (st->done) return;
    const double alpha = st->alpha;
    const int nblk = Nc + (fi ? Ng : 0);
    double rz = 0.0, xq = 0.0;
    // 8 lanes per diagonal block (6 or 4 active): lane a owns row a; the block's residual travels by width-8 shuffles
    const int sub = threadIdx.x & 7;
    const int nblk8 = (nblk + 15) / 16 * 16;   // whole warps stay together for the shuffles
    for (int blk = (blockIdx.x * blockDim.x + threadIdx.x) >> 3; blk < nblk8; blk += (gridDim.x * blockDim.x) >> 3) {
        const bool valid = blk < nblk;
        const bool cam = blk < Nc;
        const int n = cam ? 6 : 4;
        const int off = cam ? 6 * blk : 6 * Nc + 4 * (blk - Nc);
        const bool act = valid && sub < n;
        double rra = 0.0;
        if (act) {
            const double xv = x[off + sub] + alpha * p[off + sub];
            rra = r[off + sub] - alpha * q[off + sub];
            x[off + sub] = xv; r[off + sub] = rra;
            xq += xv * (b[off + sub] + rra);
        }
        double zz = 0.0;
        const double* M = cam ? Minv + 36 * size_t(blk) + n * sub : Minv + 36 * size_t(Nc) + 16 * size_t(blk - Nc) + n * sub;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const double rk = __shfl_sync(0xffffffffu, rra, k, 8);
            if (act && k < n) zz += M[k] * rk;
        }
        if (act) { z[off + sub] = zz; rz += rra * zz; }
    }
    rz = block_sum(rz, sred);
    if (threadIdx.x == 0) partial[blockIdx.x] = rz;
    xq = block_sum(xq, sred);
    if (threadIdx.x == 0) {
        partial[gridDim.x + blockIdx.x] = xq;
        __threadfence();
        const unsigned int tk = atomicAdd(&st->ticket_r, 1u);
        last = (tk == gridDim.x - 1);
    }
    __syncthreads();
    double rho_new = 0.0, xb = 0.0;
    if (last) {   // fixed-order sum of the block partials by the whole CTA (strided per thread, then the block tree)
        __threadfence();
        for (int i = threadIdx.x; i < int(gridDim.x); i += blockDim.x) { rho_new += __ldcg(partial + i); xb += __ldcg(partial + gridDim.x + i); }
        rho_new = block_sum(rho_new, sred);
        xb = block_sum(xb, sred);
    }
    if (last && threadIdx.x == 0) {
        const int k = st->k + 1;
        const double rel = sqrt(fmax(rho_new, 0.0) / st->rho0);
        int done = 0;
        if (pp.forcing) {   // Ceres' Q-criterion: Q = -1/2 x.(b + r); stop when k (Q1 - Q0)/Q1 < eta
            const double Q1 = -0.5 * xb;
            const double zeta = k * (Q1 - st->Q0) / Q1;
            st->Q0 = Q1;
            if (zeta < pp.eta) done = 1;
        }
        if (rel <= pp.rel_tolerance) done = 1;
        if (!done && k >= pp.max_iterations) done = 2;
        if (!isfinite(rho_new)) done = 3;
        st->beta = rho_new / st->rho;
        st->rho = rho_new; st->rel = rel; st->k = k; st->done = done;
        st->ticket_r = 0;
    }
}

__global__ void __launch_bounds__(PCG_BLOCK)
k_pcg_p(int nf, const double* __restrict__ z, double* __restrict__ p, const PcgState2* __restrict__ st) {
    if (st->done) return;   // converged in this iteration: p is not needed any more
    const double beta = st->beta;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < nf; j += gridDim.x * blockDim.x) p[j] = z[j] + beta * p[j];
}

// PCG start: b = reduced rhs (already reduced into `b`), x = 0, r = b, z = Minv b, p = z, rho = b.z (single CTA, once per LM step).
__global__ void __launch_bounds__(1024)
k_pcg_init2(int nf, int Nc, const double* __restrict__ Minv, const double* __restrict__ b, double* __restrict__ x,
            double* __restrict__ r, double* __restrict__ z, double* __restrict__ p, PcgState2* __restrict__ st) {
    __shared__ double sred[32];
    double acc = 0.0;
    for (int j = threadIdx.x; j < nf; j += blockDim.x) {
        x[j] = 0.0;
        const double zj = apply_minv(j, Nc, Minv, b);
        r[j] = b[j]; z[j] = zj; p[j] = zj;
        acc += b[j] * zj;
    }
    acc = block_sum(acc, sred);
    if (threadIdx.x == 0) {
        st->rho = acc; st->rho0 = acc; st->rel = acc > 0.0 ? 1.0 : 0.0; st->Q0 = 0.0; st->alpha = 0.0; st->beta = 0.0; st->k = 0;
        st->done = (acc > 0.0 && isfinite(acc)) ? 0 : (acc == 0.0 ? 1 : 3);
        st->ticket_q = 0; st->ticket_r = 0;
    }
}

}  // namespace bak
