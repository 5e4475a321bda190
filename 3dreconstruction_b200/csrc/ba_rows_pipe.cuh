// ba_rows_pipe.cuh — K4/K5a as a PERSISTENT, software-pipelined version of k_schur_rows (scatter mode 5).
//
// Same arithmetic, same static schedule and same partial-row output as ba_rows.cuh (so K5b k_camera_gather is unchanged and the
// result is bit-identical to mode 4); what changes is how the bytes arrive.  ncu on the one-CTA-per-tile kernel shows ~36 % of its
// stall samples on a chain of three dependent memory latencies per CTA (tile record -> schedule record + J -> x[camera]) and
// nothing in flight while a CTA is in its later phases.  Here a CTA walks tiles t, t+G, t+2G, ... and every thread copies ITS OWN
// ten 16-byte pieces of the NEXT tile (9 Jacobian column pairs + the schedule record) into its own shared-memory slots with
// cp.async (LDGSTS.128, L2-only) while the CTA computes the current tile from registers; tile records are prefetched two tiles
// ahead and the next tile's Hinv lines are touched with prefetch.global.L2.  No thread ever reads another thread's landing slot,
// so the pipeline needs no barrier of its own: cp.async.wait_group orders a thread's copies before its reads.
// Shared memory: 40 KB landing + 24.6 KB exchange buffers per CTA => 3 CTAs/SM, ~120 KB of loads in flight per SM at all times.
#pragma once
#include "ba_rows.cuh"

namespace bak {

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

template <int MODE, bool FI>
struct RowsPipeCfg {
    static constexpr int NJ = FI ? 13 : 9;
    static constexpr int NROW = NJ + (MODE == SWEEP_RHS ? 1 : 0);                      // + the residual for K4
    static constexpr size_t SMEM = size_t(NROW) * TILE_OBS * sizeof(double2) + size_t(TILE_OBS) * sizeof(int4);
    static constexpr int CTAS_PER_SM = FI ? 2 : 3;
};

// One point with more than TILE_OBS observations (rare): two passes straight from global memory, every observation is its own row.
template <int MODE, bool FI>
__device__ __noinline__ void rows_long_tile(const View& V, const RowSched& S, const double* __restrict__ xin, double* __restrict__ atomic_out,
                                            int o0, int n, int hbase, int p, double (*sV)[3], double* sred, double* yi) {
    const size_t ld = V.ldo;
    const int ioff = 6 * V.Nc;
    const int o1 = o0 + n;
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
        double2* dst = reinterpret_cast<double2*>(S.rows + size_t(ROW_DOUBLES) * (hbase + (o - o0)));
        __stcg(dst, make_double2(jc[0].x * tt.x + jc[0].y * tt.y, jc[1].x * tt.x + jc[1].y * tt.y));
        __stcg(dst + 1, make_double2(jc[2].x * tt.x + jc[2].y * tt.y, jc[3].x * tt.x + jc[3].y * tt.y));
        __stcg(dst + 2, make_double2(jc[4].x * tt.x + jc[4].y * tt.y, jc[5].x * tt.x + jc[5].y * tt.y));
        if (FI) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const double contrib = ji[k].x * tt.x + ji[k].y * tt.y;
                if (V.Ng > 1) red_add(atomic_out + ioff + 4 * g + k, contrib); else yi[k] += contrib;
            }
        }
    }
}

template <int MODE, bool FI>
__global__ void __launch_bounds__(TILE_OBS, RowsPipeCfg<MODE, FI>::CTAS_PER_SM)
k_schur_rows_pipe(View V, RowSched S, const double* __restrict__ xin, double* __restrict__ atomic_out, const int* __restrict__ done_flag) {
    using Cfg = RowsPipeCfg<MODE, FI>;
    constexpr int NJ = Cfg::NJ, NROW = Cfg::NROW;
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    double2* sL = reinterpret_cast<double2*>(dyn_smem) + threadIdx.x;                          // [NROW][TILE_OBS], this thread's column
    int4* sT = reinterpret_cast<int4*>(reinterpret_cast<double2*>(dyn_smem) + size_t(NROW) * TILE_OBS) + threadIdx.x;
    __shared__ double sW[TILE_OBS][3];
    __shared__ double sV[TILE_OBS][3];
    __shared__ __align__(16) double sC[TILE_OBS][6];
    __shared__ double sred[32];
    if (done_flag && *done_flag) return;
    const int G = gridDim.x, nT = V.nTiles;
    int t = blockIdx.x;
    if (t >= nT) return;
    const size_t ld = V.ldo;
    const int ioff = 6 * V.Nc;
    const int4 zero4 = make_int4(0, 0, 0, 0);

    // this thread's copies of tile `ti`: 9 (13) Jacobian column pairs [+ residual] + the schedule record
    auto issue = [&](const int4 ti) {
        if (ti.y <= TILE_OBS && int(threadIdx.x) < ti.y) {
            const int o = ti.x + threadIdx.x;
            const double2* J = V.J2 + o;
#pragma unroll
            for (int k = 0; k < NJ; ++k) cp_async16(sL + k * TILE_OBS, J + k * ld);
            if (MODE == SWEEP_RHS) cp_async16(sL + NJ * TILE_OBS, V.r2 + o);
            cp_async16(sT, S.tmeta + o);
        }
        cp_async_commit();
    };

    int4 ti = __ldg(S.tile_info + t);
    int4 tn = (t + G < nT) ? __ldg(S.tile_info + t + G) : zero4;
    issue(ti);

    for (; t < nT; t += G) {
        const int4 tnn = (t + 2 * G < nT) ? __ldg(S.tile_info + t + 2 * G) : zero4;   // tile record two tiles ahead
        const int o0 = ti.x, n = ti.y, hbase = ti.z, pfirst = ti.w;
        double yi[4] = {0, 0, 0, 0};
        if (n <= TILE_OBS) {
            const bool act = int(threadIdx.x) < n;
            int4 tm = make_int4(0, 0, -1, 0);
            double2 jc[6], jp[3], ji[4], rr = make_double2(0.0, 0.0);
            cp_async_wait_all();
            if (act) {
                tm = *sT;
#pragma unroll
                for (int k = 0; k < 6; ++k) jc[k] = sL[k * TILE_OBS];
#pragma unroll
                for (int k = 0; k < 3; ++k) jp[k] = sL[(6 + k) * TILE_OBS];
                if (FI) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) ji[k] = sL[(9 + k) * TILE_OBS];
                }
                if (MODE == SWEEP_RHS) rr = sL[NJ * TILE_OBS];
            }
            asm volatile("" ::: "memory");   // the slot reads above stay before the refills below
            if (t + G < nT) {
                issue(tn);
                if (MODE == SWEEP_MATVEC && threadIdx.x < 24) prefetch_l2(V.Hinv + 6 * size_t(tn.w) + 16 * threadIdx.x);   // next tile's Hinv lines
            }
            const int npt_obs = tm.w >> 16, pstart = tm.w & 0xffff;   // this thread as a point thread
            double2 h0 = make_double2(0, 0), h1 = h0, h2 = h0;
            if (MODE == SWEEP_MATVEC && npt_obs > 0) {
                const double2* H = reinterpret_cast<const double2*>(V.Hinv + 6 * size_t(pfirst + threadIdx.x));
                h0 = __ldg(H); h1 = __ldg(H + 1); h2 = __ldg(H + 2);
            }
            const int c = tm.x & 0xffffff;
            const int slot = tm.y & 0xff, n1 = (tm.y >> 8) & 0xff, hrank = (tm.y >> 16) & 0xff, lpt = (tm.y >> 24) & 0xff;
            const int pk = tm.z;
            const int g = (FI && act) ? V.cam_group[c] : 0;
            double2 u = make_double2(0.0, 0.0);
            double2 tt = make_double2(0.0, 0.0);
            if (MODE == SWEEP_MATVEC) {
                if (act) {
                    const double2* xc = reinterpret_cast<const double2*>(xin + 6 * c);
                    const double2 x0 = __ldg(xc), x1 = __ldg(xc + 1), x2 = __ldg(xc + 2);
                    u.x = jc[0].x * x0.x + jc[1].x * x0.y + jc[2].x * x1.x + jc[3].x * x1.y + jc[4].x * x2.x + jc[5].x * x2.y;
                    u.y = jc[0].y * x0.x + jc[1].y * x0.y + jc[2].y * x1.x + jc[3].y * x1.y + jc[4].y * x2.x + jc[5].y * x2.y;
                    if (FI) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) { const double xv = __ldg(xin + ioff + 4 * g + k); u.x += ji[k].x * xv; u.y += ji[k].y * xv; }
                    }
#pragma unroll
                    for (int a = 0; a < 3; ++a) sW[threadIdx.x][a] = jp[a].x * u.x + jp[a].y * u.y;
                }
                __syncthreads();
                if (npt_obs > 0) {
                    double W0 = 0, W1 = 0, W2 = 0;
                    for (int i = pstart, e = pstart + npt_obs; i < e; ++i) { W0 += sW[i][0]; W1 += sW[i][1]; W2 += sW[i][2]; }
                    sV[threadIdx.x][0] = h0.x * W0 + h0.y * W1 + h1.x * W2;
                    sV[threadIdx.x][1] = h0.y * W0 + h1.y * W1 + h2.x * W2;
                    sV[threadIdx.x][2] = h1.x * W0 + h2.x * W1 + h2.y * W2;
                }
                __syncthreads();
                if (act) {
                    const double* v = sV[lpt];
                    tt.x = u.x - (jp[0].x * v[0] + jp[1].x * v[1] + jp[2].x * v[2]);
                    tt.y = u.y - (jp[0].y * v[0] + jp[1].y * v[1] + jp[2].y * v[2]);
                }
            } else if (act) {   // SWEEP_RHS: t = r - E vp
                const double* v = V.vp + 3 * size_t(pfirst + lpt);
                tt.x = rr.x - (jp[0].x * v[0] + jp[1].x * v[1] + jp[2].x * v[2]);
                tt.y = rr.y - (jp[0].y * v[0] + jp[1].y * v[1] + jp[2].y * v[2]);
            }
            if (act) {
                double2* sc = reinterpret_cast<double2*>(sC[slot]);
                sc[0] = make_double2(jc[0].x * tt.x + jc[0].y * tt.y, jc[1].x * tt.x + jc[1].y * tt.y);
                sc[1] = make_double2(jc[2].x * tt.x + jc[2].y * tt.y, jc[3].x * tt.x + jc[3].y * tt.y);
                sc[2] = make_double2(jc[4].x * tt.x + jc[4].y * tt.y, jc[5].x * tt.x + jc[5].y * tt.y);
                if (FI) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) yi[k] = ji[k].x * tt.x + ji[k].y * tt.y;
                    if (V.Ng > 1) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) red_add(atomic_out + ioff + 4 * g + k, yi[k]);
                    }
                }
            }
            __syncthreads();
            if (n1 > 1) {   // level 1: chunk leaders fold their <= L1_CHUNK rows
                double2* row = reinterpret_cast<double2*>(sC[threadIdx.x]);
                double2 a = row[0], bq = row[1], cq = row[2];
                for (int i = 1; i < n1; ++i) {
                    const double2 a2 = row[3 * i], b2 = row[3 * i + 1], c2 = row[3 * i + 2];
                    a.x += a2.x; a.y += a2.y; bq.x += b2.x; bq.y += b2.y; cq.x += c2.x; cq.y += c2.y;
                }
                row[0] = a; row[1] = bq; row[2] = cq;
            }
            __syncthreads();
            if (pk >= 0) {   // level 2: segment heads add their chunk leaders and store one 64-byte row
                const int len = pk >> 22;
                const int j = threadIdx.x, end = j + len;
                const double2* row = reinterpret_cast<const double2*>(sC[j]);
                double2 a = row[0], bq = row[1], cq = row[2];
                for (int q = (j / L1_CHUNK + 1) * L1_CHUNK; q < end; q += L1_CHUNK) {
                    const double2* r2 = reinterpret_cast<const double2*>(sC[q]);
                    const double2 a2 = r2[0], b2 = r2[1], c2 = r2[2];
                    a.x += a2.x; a.y += a2.y; bq.x += b2.x; bq.y += b2.y; cq.x += c2.x; cq.y += c2.y;
                }
                double2* dst = reinterpret_cast<double2*>(S.rows + size_t(ROW_DOUBLES) * (hbase + hrank));
                __stcg(dst, a); __stcg(dst + 1, bq); __stcg(dst + 2, cq);
            }
            // no barrier needed before the next tile: its first sC write comes after two more barriers, its first sW write
            // after this tile's last sW read (barrier 2) — except on the RHS path, which has no barriers 1-2:
            if (MODE == SWEEP_RHS || FI) __syncthreads();
        } else {
            if (t + G < nT) issue(tn);
            __syncthreads();
            rows_long_tile<MODE, FI>(V, S, xin, atomic_out, o0, n, hbase, pfirst, sV, sred, yi);
            __syncthreads();
        }
        if (FI && V.Ng == 1) {   // one shared intrinsics group: per-tile partial, summed in tile order by k_camera_gather
#pragma unroll
            for (int k = 0; k < 4; ++k) { const double s = block_sum(yi[k], sred); if (threadIdx.x == 0) S.tile_intr[4 * size_t(t) + k] = s; }
            __syncthreads();
        }
        ti = tn; tn = tnn;
    }
}

}  // namespace bak
