// ba_pipe.cuh — K4/K5 as a persistent, warp-specialised kernel: one producer warp streams each tile's Jacobian rows
// from HBM into a shared-memory ring with 1-D bulk async copies (cp.async.bulk -> SASS UBLKCP) completing on mbarriers,
// eight consumer warps do the math; the camera-side scatter is the atomic-free scheme of ba_scatter.cuh.  
//
// Why a pipeline: the sweep has four phases per tile separated by barriers (load, per-point reduce, scatter to shared
// memory, segment-head read-modify-write).  With plain loads a CTA has nothing in flight while it is in the later phases,
// and only two CTAs fit per SM next to their private accumulators; decoupling the loads keeps >= 2 tiles (74 KB) per
// CTA in flight regardless of what the consumers are doing.
#pragma once
#include "ba_scatter.cuh"

namespace bak {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
// global -> shared bulk copy, `bytes` a multiple of 16, both addresses 16-byte aligned; completes on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bar_consumers() { asm volatile("bar.sync 1, %0;" ::"n"(TILE_OBS) : "memory"); }

template <int MODE, bool FI>
struct PipeCfg {
    static constexpr int NROW = (FI ? 13 : 9) + (MODE == SWEEP_RHS ? 1 : 0);   // J column pairs (+ the residual row for K4)
    static constexpr int STAGES = 2;
    static constexpr int THREADS = TILE_OBS + 32;
    static constexpr size_t SMEM = size_t(STAGES) * NROW * TILE_OBS * sizeof(double2)   // landing ring
                                   + size_t(TILE_OBS) * (3 + 3 + 6) * sizeof(double)    // sW, sV, sC
                                   + 2 * STAGES * sizeof(uint64_t) + 32 * sizeof(double);
    static constexpr int CTAS_PER_SM = (2 * (SMEM + 1024) <= 227 * 1024) ? 2 : 1;
};

template <int MODE, bool FI>
__global__ void __launch_bounds__(TILE_OBS + 32, PipeCfg<MODE, FI>::CTAS_PER_SM)
k_schur_pipe(View V, ScatterSched Sch, const double* __restrict__ xin, const int* __restrict__ done_flag) {
    using Cfg = PipeCfg<MODE, FI>;
    constexpr int NROW = Cfg::NROW, STAGES = Cfg::STAGES, NJ = FI ? 13 : 9;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double2* sJ = reinterpret_cast<double2*>(smem_raw);                                   // [STAGES][NROW][TILE_OBS]
    double* sW = reinterpret_cast<double*>(sJ + size_t(STAGES) * NROW * TILE_OBS);        // [TILE_OBS][3]
    double* sV = sW + 3 * TILE_OBS;                                                       // [TILE_OBS][3]
    double* sC = sV + 3 * TILE_OBS;                                                       // [TILE_OBS][6]
    double* sred = sC + 6 * TILE_OBS;                                                     // [32]
    uint64_t* full = reinterpret_cast<uint64_t*>(sred + 32);                              // [STAGES]
    uint64_t* empty = full + STAGES;                                                      // [STAGES]

    if (done_flag && *done_flag) return;
    const int G = gridDim.x, b = blockIdx.x;
    const int t0 = int((long long)V.nTiles * b / G), t1 = int((long long)V.nTiles * (b + 1) / G);
    double* acc = Sch.scratch + size_t(b) * Sch.ldf;
    const size_t ld = V.ldo;
    const int ioff = 6 * V.Nc;

    // prologue: clear this CTA's private accumulator, arm the barriers
    for (int j = threadIdx.x; j < Sch.ldf; j += blockDim.x) __stcg(acc + j, 0.0);
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (threadIdx.x >= TILE_OBS) {
        // ================= producer warp: one lane issues the bulk copies =================
        if (threadIdx.x == TILE_OBS) {
            int it = 0;
            for (int t = t0; t < t1; ++t) {
                const int o0 = V.tile_off[t], n = V.tile_off[t + 1] - o0;
                if (n > TILE_OBS) continue;                         // long-point tile: consumers read it straight from global
                const int s = it % STAGES;
                mbar_wait(empty + s, ((it / STAGES) & 1) ^ 1);      // slot free (passes immediately the first time round)
                const uint32_t bytes = uint32_t(n) * sizeof(double2);
                mbar_expect_tx(full + s, bytes * NROW);
                double2* dst = sJ + size_t(s) * NROW * TILE_OBS;
#pragma unroll
                for (int k = 0; k < NJ; ++k) bulk_g2s(dst + k * TILE_OBS, V.J2 + k * ld + o0, bytes, full + s);
                if (MODE == SWEEP_RHS) bulk_g2s(dst + NJ * TILE_OBS, V.r2 + o0, bytes, full + s);
                ++it;
            }
        }
        return;
    }

    // ================= consumer warps =================
    double yi[4] = {0, 0, 0, 0};
    int it = 0;
    for (int t = t0; t < t1; ++t) {
        const int o0 = V.tile_off[t], o1 = V.tile_off[t + 1];
        if (o1 - o0 <= TILE_OBS) {
            const int o = o0 + threadIdx.x;
            const bool act = o < o1;
            int c = 0, p = 0, g = 0, slot = 0, pk = -1;
            if (act) { c = V.obs_cam[o]; p = V.obs_pt[o]; slot = Sch.slot[o]; pk = Sch.spack[o]; if (FI) g = V.cam_group[c]; }
            const int pfirst = V.obs_pt[o0], plast = V.obs_pt[o1 - 1];
            const int s = it % STAGES;
            const double2* J = sJ + size_t(s) * NROW * TILE_OBS + threadIdx.x;
            mbar_wait(full + s, (it / STAGES) & 1);
            double2 jc[6], jp[3], ji[4], rr = make_double2(0.0, 0.0);
            double2 u = make_double2(0.0, 0.0);
            if (act) {
#pragma unroll
                for (int k = 0; k < 6; ++k) jc[k] = J[k * TILE_OBS];
#pragma unroll
                for (int k = 0; k < 3; ++k) jp[k] = J[(6 + k) * TILE_OBS];
                if (FI) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) ji[k] = J[(9 + k) * TILE_OBS];
                }
                if (MODE == SWEEP_RHS) rr = J[NJ * TILE_OBS];
                if (MODE == SWEEP_MATVEC) {
                    const double* xc = xin + 6 * c;
#pragma unroll
                    for (int k = 0; k < 6; ++k) { const double xv = __ldg(xc + k); u.x += jc[k].x * xv; u.y += jc[k].y * xv; }
                    if (FI) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) { const double xv = __ldg(xin + ioff + 4 * g + k); u.x += ji[k].x * xv; u.y += ji[k].y * xv; }
                    }
#pragma unroll
                    for (int a = 0; a < 3; ++a) sW[3 * threadIdx.x + a] = jp[a].x * u.x + jp[a].y * u.y;
                }
            }
            bar_consumers();                                         // every consumer has its rows in registers (and sW is written)
            if (threadIdx.x == 0) mbar_arrive(empty + s);            // hand the landing slot back to the producer
            ++it;
            double2 tt = make_double2(0.0, 0.0);
            if (MODE == SWEEP_MATVEC) {
                const int pj = pfirst + threadIdx.x;
                if (pj <= plast) {
                    double W0 = 0, W1 = 0, W2 = 0;
                    for (int i = V.pt_off[pj] - o0, e = V.pt_off[pj + 1] - o0; i < e; ++i) { W0 += sW[3 * i]; W1 += sW[3 * i + 1]; W2 += sW[3 * i + 2]; }
                    const double* H = V.Hinv + 6 * size_t(pj);
                    sV[3 * threadIdx.x] = H[0] * W0 + H[1] * W1 + H[2] * W2;
                    sV[3 * threadIdx.x + 1] = H[1] * W0 + H[3] * W1 + H[4] * W2;
                    sV[3 * threadIdx.x + 2] = H[2] * W0 + H[4] * W1 + H[5] * W2;
                }
                bar_consumers();
                if (act) {
                    const double* v = sV + 3 * (p - pfirst);
                    tt.x = u.x - (jp[0].x * v[0] + jp[1].x * v[1] + jp[2].x * v[2]);
                    tt.y = u.y - (jp[0].y * v[0] + jp[1].y * v[1] + jp[2].y * v[2]);
                }
            } else if (act) {   // SWEEP_RHS: t = r - E vp
                const double* v = V.vp + 3 * size_t(p);
                tt.x = rr.x - (jp[0].x * v[0] + jp[1].x * v[1] + jp[2].x * v[2]);
                tt.y = rr.y - (jp[0].y * v[0] + jp[1].y * v[1] + jp[2].y * v[2]);
            }
            if (act) {
                double2* sc = reinterpret_cast<double2*>(sC + 6 * slot);
                sc[0] = make_double2(jc[0].x * tt.x + jc[0].y * tt.y, jc[1].x * tt.x + jc[1].y * tt.y);
                sc[1] = make_double2(jc[2].x * tt.x + jc[2].y * tt.y, jc[3].x * tt.x + jc[3].y * tt.y);
                sc[2] = make_double2(jc[4].x * tt.x + jc[4].y * tt.y, jc[5].x * tt.x + jc[5].y * tt.y);
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
            bar_consumers();
            if (pk >= 0) {   // this thread sits at a segment head of the tile's camera-sorted order
                const int cam = pk & 0x3fffff, len = pk >> 22;
                double s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0, s5 = 0;
                const double2* row = reinterpret_cast<const double2*>(sC + 6 * threadIdx.x);
                for (int i = 0; i < len; ++i) {
                    const double2 a = row[3 * i], bq = row[3 * i + 1], cq = row[3 * i + 2];
                    s0 += a.x; s1 += a.y; s2 += bq.x; s3 += bq.y; s4 += cq.x; s5 += cq.y;
                }
                double2* dst = reinterpret_cast<double2*>(acc + 6 * size_t(cam));
                double2 a = __ldcg(dst), bq = __ldcg(dst + 1), cq = __ldcg(dst + 2);
                a.x += s0; a.y += s1; bq.x += s2; bq.y += s3; cq.x += s4; cq.y += s5;
                __stcg(dst, a); __stcg(dst + 1, bq); __stcg(dst + 2, cq);
            }
            // no barrier here: the next tile's first write to sC comes after two more bar_consumers()
        } else {
            // ---- one point with more than TILE_OBS observations (rare): two passes straight from global memory ----
            bar_consumers();
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
                for (int a = 0; a < 3; ++a) {   // consumer-only block sum (the producer warp must not be in the barrier)
                    double w = warp_sum(W[a]);
                    if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = w;
                    bar_consumers();
                    double tot = 0.0;
                    for (int q = 0; q < TILE_OBS / 32; ++q) tot += sred[q];
                    W[a] = tot;
                    bar_consumers();
                }
                const double* H = V.Hinv + 6 * size_t(p);
                v[0] = H[0] * W[0] + H[1] * W[1] + H[2] * W[2]; v[1] = H[1] * W[0] + H[3] * W[1] + H[4] * W[2]; v[2] = H[2] * W[0] + H[4] * W[1] + H[5] * W[2];
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
            bar_consumers();
        }
    }
    if (FI && V.Ng == 1) {   // shared intrinsics group: consumer-wide sum of the per-thread partials, one add into the private copy
        bar_consumers();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const double w = warp_sum(yi[k]);
            if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = w;
            bar_consumers();
            if (threadIdx.x == 0) { double tot = 0.0; for (int q = 0; q < TILE_OBS / 32; ++q) tot += sred[q]; __stcg(acc + ioff + k, __ldcg(acc + ioff + k) + tot); }
            bar_consumers();
        }
    }
}

// Stage 1 of the copy reduction: part[r][j] = sum over the r-th slice of the G private copies (fixed order).
__global__ void __launch_bounds__(256)
k_reduce_copies(int nf, int G, int ldf, const double* __restrict__ scratch, int R, double* __restrict__ part, const int* __restrict__ done_flag) {
    if (done_flag && *done_flag) return;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nf) return;
    const int r = blockIdx.y;
    const int g0 = int((long long)G * r / R), g1 = int((long long)G * (r + 1) / R);
    double s = 0.0;
    int g = g0;
    for (; g + 4 <= g1; g += 4) {
        const double a0 = __ldcg(scratch + size_t(g) * ldf + j), a1 = __ldcg(scratch + size_t(g + 1) * ldf + j);
        const double a2 = __ldcg(scratch + size_t(g + 2) * ldf + j), a3 = __ldcg(scratch + size_t(g + 3) * ldf + j);
        s += a0; s += a1; s += a2; s += a3;
    }
    for (; g < g1; ++g) s += __ldcg(scratch + size_t(g) * ldf + j);
    part[size_t(r) * ldf + j] = s;
}

}  // namespace bak
