// ba_pcg_cluster.cuh — everything a PCG iteration does after the Schur product, in ONE thread-block cluster.
//
// After K5a/K5b an iteration still has three small vector steps separated by global scalars:
//     q = y + Df2 p,  alpha = rho / p.q   |   x += alpha p, r -= alpha q, z = Minv r, rho' = r.z, beta, stop test   |   p = z + beta p
// On 10 668 unknowns (Venice) each step is a few microseconds of work, but as separate kernels with last-block tickets they cost
// 8-10 us each plus launch gaps (k_pcg_q[_p2p], k_pcg_xr, k_pcg_p: ~28 us per iteration; at 8 GPUs that is as long as the local K5a).
// Here one cluster of TAIL_CL CTAs (one launch) owns the whole reduced vector: every CTA owns a contiguous range of diagonal blocks,
// the three scalars are reduced through distributed shared memory (every CTA writes its partial into its own shared memory, one
// cluster barrier, every CTA adds the TAIL_CL partials in rank order: no global memory, no atomics, identical bits everywhere).
//
// Source of y (mode): 0 = the gather already produced q = y + Df2 p and one share of p.q per gather CTA (dotp[0..ndot), single GPU):
//                         every CTA adds the shares in index order (same bits everywhere, no cluster reduction, no second pass over y);
//                     1 = y is in local memory, already all-reduced (NCCL path, atomics scatter);
//                     2 = pull the W ranks' slots from peer memory and add them in rank order (ba_p2p.cuh), after waiting for the flags.
#pragma once
#include <cooperative_groups.h>

#include "ba_p2p.cuh"

namespace bak {

#ifndef BA_TAIL_PREFETCH
#define BA_TAIL_PREFETCH 1
#endif
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
constexpr int TAIL_CL = 8;          // CTAs in the cluster (portable maximum)
constexpr int TAIL_CL_WIDE = 16;    // non-portable cluster size (cudaFuncAttributeNonPortableClusterSizeAllowed): large reduced systems, when the device can place it
constexpr int TAIL_THREADS = 1024;

struct TailArgs {
    int Nc, Ng, fi, mode;
    const double* Minv; const double* b; const double* Df2;
    double* y; double* p; double* x; double* r; double* z;
    const double* dotp; int ndot;        // mode 0: the gather CTAs' shares of p.q
    const double* G; double* vt;         // pose-factored matvec (ba_rows.cuh: k_pose_factors): vt = [G_c p_c | p of the intrinsics entries], or null
    PcgState2* st;
    PcgParams pp;
    PeerView P;
    unsigned int seq;
};

// Sum of one double per CTA over the cluster, the same value (bit for bit) in every thread of every CTA.  `slot` is this CTA's own
// shared-memory cell for THIS reduction (a different cell per reduction, so no second barrier is needed before reuse).
template <int CL>
__device__ __forceinline__ double cluster_total(cooperative_groups::cluster_group& cl, double v, double* slot, double* sred) {
    v = block_sum(v, sred);
    if (threadIdx.x == 0) *slot = v;
    cl.sync();
    double tot = 0.0;
#pragma unroll
    for (int k = 0; k < CL; ++k) tot += *cl.map_shared_rank(slot, k);
    return tot;
}

template <int CL>
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(TAIL_THREADS) k_pcg_tail(TailArgs A) {
    namespace cg = cooperative_groups;
    __shared__ double sred[32];
    __shared__ double cell[3];
    cg::cluster_group cl = cg::this_cluster();
    PcgState2* st = A.st;
    pdl_trigger();
    const int Nc = A.Nc, nblk = Nc + (A.fi ? A.Ng : 0);
    const int per = ((nblk + CL - 1) / CL + 3) & ~3;          // diagonal blocks per CTA, whole warps of 8-lane groups
    const int b0 = int(cl.block_rank()) * per, b1 = min(nblk, b0 + per);
    const int sub = threadIdx.x & 7, grp = threadIdx.x >> 3;
    constexpr int NGRP = TAIL_THREADS / 8;
#if BA_TAIL_PREFETCH
    // Everything this CTA reads except the gather's output (y, dotp) was last written by the PREVIOUS tail or is static during the solve:
    // pull those lines into this SM's L1 while the predecessor grid is still running, so that the dependent load phases after the wait
    // (x, r, p, b, Df2 -> Minv rows -> z, p, G rows) are L1 hits.  The kernel is a chain of latencies, not of bytes.
    for (int base = b0; base < b1; base += NGRP) {
        const int blk = base + grp;
        if (blk < b1) {
            const bool cam = blk < Nc;
            const int n = cam ? 6 : 4, off = cam ? 6 * blk : 6 * Nc + 4 * (blk - Nc);
            if (sub < n) {
                prefetch_l1(A.x + off + sub); prefetch_l1(A.r + off + sub); prefetch_l1(A.p + off + sub); prefetch_l1(A.b + off + sub);
                if (A.mode != 0) prefetch_l1(A.Df2 + off + sub);
                const double* M = cam ? A.Minv + 36 * size_t(blk) + n * sub : A.Minv + 36 * size_t(Nc) + 16 * size_t(blk - Nc) + n * sub;
                prefetch_l1(M); prefetch_l1(M + n - 1);
                if (A.vt && cam) { const double* Gr = A.G + 36 * size_t(blk) + 6 * sub; prefetch_l1(Gr); prefetch_l1(Gr + 5); }
            }
        }
    }
#endif
    pdl_wait();
    if (st->done) return;                        // uniform over the cluster: nobody reaches a barrier
    const double rho = st->rho;
    double alpha = st->alpha;

    // ---- q = y + Df2 p, p.q, alpha ------------------------------------------------------------------------------------------
    if (A.mode != 0) {
        size_t off = 0;
        if (A.mode == 2) {
            // this rank's slot is complete (the gather kernel has finished): one system-scope fence, then tell every rank
            if (cl.block_rank() == 0 && threadIdx.x < A.P.W) { __threadfence_system(); st_release_sys(A.P.flags[threadIdx.x] + A.P.rank, A.seq); }
            if (threadIdx.x < A.P.W) {
                const unsigned int* f = A.P.flags[A.P.rank] + threadIdx.x;
                const long long t0 = clock64();
                while (int(ld_acquire_sys(f) - A.seq) < 0) {
                    if ((*reinterpret_cast<volatile int*>(A.P.timeout_flag) & 8) || clock64() - t0 > 4000000000LL) { atomicOr(A.P.timeout_flag, 8); break; }
                }
            }
            __syncthreads();
            off = size_t(A.seq & 1u) * A.P.ldf;
        }
        double acc = 0.0;
        if (A.mode == 2) {
            // peer loads cost an NVLink round trip each: issue them for UNR blocks x W ranks at once, then add in rank order
            constexpr int UNR = 2;
            for (int base = b0 + grp; base < b1; base += UNR * NGRP) {
                double v[UNR][P2P_MAX_RANKS];
                int jj[UNR];
#pragma unroll
                for (int u = 0; u < UNR; ++u) {
                    const int blk = base + u * NGRP;
                    const bool cam = blk < Nc;
                    const int n = cam ? 6 : 4;
                    jj[u] = (blk < b1 && sub < n) ? (cam ? 6 * blk : 6 * Nc + 4 * (blk - Nc)) + sub : -1;
#pragma unroll
                    for (int rk = 0; rk < P2P_MAX_RANKS; ++rk) v[u][rk] = (jj[u] >= 0 && rk < A.P.W) ? ld_relaxed_sys_f64(A.P.data[rk] + off + jj[u]) : 0.0;
                }
#pragma unroll
                for (int u = 0; u < UNR; ++u) {
                    if (jj[u] < 0) continue;
                    double s = 0.0;
#pragma unroll
                    for (int rk = 0; rk < P2P_MAX_RANKS; ++rk) if (rk < A.P.W) s += v[u][rk];
                    const double pj = A.p[jj[u]];
                    s += A.Df2[jj[u]] * pj;
                    acc += pj * s;
                    A.y[jj[u]] = s;
                }
            }
        } else {
            for (int blk = b0 + grp; blk < b1; blk += NGRP) {
                const bool cam = blk < Nc;
                const int n = cam ? 6 : 4, j = (cam ? 6 * blk : 6 * Nc + 4 * (blk - Nc)) + sub;
                if (sub < n) {
                    const double pj = A.p[j];
                    const double s = A.y[j] + A.Df2[j] * pj;
                    acc += pj * s;
                    A.y[j] = s;
                }
            }
        }
        const double pq = cluster_total<CL>(cl, acc, &cell[0], sred);
        if (!(pq > 0.0) || !isfinite(pq)) {      // breakdown: every CTA sees the same pq
            if (cl.block_rank() == 0 && threadIdx.x == 0) { st->done = 3; st->alpha = 0.0; }
            cl.sync();                           // nobody leaves while its shared memory may still be read
            return;
        }
        alpha = rho / pq;
    } else {
        __shared__ double s_pq;
        double acc = 0.0;
        for (int i = threadIdx.x; i < A.ndot; i += TAIL_THREADS) acc += A.dotp[i];
        acc = block_sum(acc, sred);
        if (threadIdx.x == 0) s_pq = acc;
        __syncthreads();
        const double pq = s_pq;
        if (!(pq > 0.0) || !isfinite(pq)) {      // breakdown: every CTA sees the same pq; no cluster barrier has been entered yet
            if (cl.block_rank() == 0 && threadIdx.x == 0) { st->done = 3; st->alpha = 0.0; }
            return;
        }
        alpha = rho / pq;
    }

    // ---- x += alpha p, r -= alpha q, z = Minv r; r.z and x.(b + r) ----------------------------------------------------------------
    double rz = 0.0, xq = 0.0;
    for (int base = b0; base < b1; base += NGRP) {   // whole warps iterate together (width-8 shuffles below)
        const int blk = base + grp;
        const bool valid = blk < b1;
        const bool cam = blk < Nc;
        const int n = cam ? 6 : 4;
        const int off = cam ? 6 * blk : 6 * Nc + 4 * (blk - Nc);
        const bool act = valid && sub < n;
        double rra = 0.0;
        if (act) {
            const double xv = A.x[off + sub] + alpha * A.p[off + sub];
            rra = A.r[off + sub] - alpha * A.y[off + sub];
            A.x[off + sub] = xv; A.r[off + sub] = rra;
            xq += xv * (A.b[off + sub] + rra);
        }
        double zz = 0.0;
        const double* M = valid ? (cam ? A.Minv + 36 * size_t(blk) + n * sub : A.Minv + 36 * size_t(Nc) + 16 * size_t(blk - Nc) + n * sub) : A.Minv;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const double rk = __shfl_sync(0xffffffffu, rra, k, 8);
            if (act && k < n) zz += M[k] * rk;
        }
        if (act) { A.z[off + sub] = zz; rz += rra * zz; }
    }
    // r.z and x.(b + r) in ONE cluster barrier
    rz = block_sum(rz, sred);
    xq = block_sum(xq, sred);
    if (threadIdx.x == 0) { cell[1] = rz; cell[2] = xq; }
    cl.sync();
    double rho_new = 0.0, xb = 0.0;
#pragma unroll
    for (int k = 0; k < CL; ++k) { const double* c = cl.map_shared_rank(&cell[0], k); rho_new += c[1]; xb += c[2]; }

    // ---- stop rules (identical in every thread; CTA 0 records them), then p = z + beta p -----------------------------------------
    const int k = st->k + 1;
    const double rel = sqrt(fmax(rho_new, 0.0) / st->rho0);
    int done = 0;
    double Q1 = 0.0;
    if (A.pp.forcing) {   // Ceres' Q-criterion: Q = -1/2 x.(b + r); stop when k (Q1 - Q0)/Q1 < eta
        Q1 = -0.5 * xb;
        const double zeta = k * (Q1 - st->Q0) / Q1;
        if (zeta < A.pp.eta) done = 1;
    }
    if (rel <= A.pp.rel_tolerance) done = 1;
    if (!done && k >= A.pp.max_iterations) done = 2;
    if (!isfinite(rho_new)) done = 3;
    const double beta = rho_new / rho;
    if (!done) {
        for (int base = b0; base < b1; base += NGRP) {   // whole warps iterate together (width-8 shuffles below)
            const int blk = base + grp;
            const bool valid = blk < b1, cam = blk < Nc;
            const int n = cam ? 6 : 4, off = cam ? 6 * blk : 6 * Nc + 4 * (blk - Nc);
            const bool act = valid && sub < n;
            double pn = 0.0;
            if (act) { pn = A.z[off + sub] + beta * A.p[off + sub]; A.p[off + sub] = pn; }
            if (A.vt) {   // what the next pose-factored K5a gathers instead of p: row `sub` of G_c times the new p_c
                const double* Gr = (act && cam) ? A.G + 36 * size_t(blk) + 6 * sub : nullptr;
                double vv = 0.0;
#pragma unroll
                for (int k = 0; k < 6; ++k) {
                    const double pk = __shfl_sync(0xffffffffu, pn, k, 8);
                    if (Gr) vv += Gr[k] * pk;
                }
                if (act) A.vt[cam ? VT_STRIDE * size_t(blk) + sub : off + (VT_STRIDE - 6) * Nc + sub] = cam ? vv : pn;
            }
        }
    }
    cl.sync();            // every CTA has read st->k, st->rho0, st->Q0 and every partial before anything is overwritten or freed
    if (cl.block_rank() == 0 && threadIdx.x == 0) {
        if (A.pp.forcing) st->Q0 = Q1;
        st->alpha = alpha; st->beta = beta; st->rho = rho_new; st->rel = rel; st->k = k; st->done = done;
    }
}

}  // namespace bak
