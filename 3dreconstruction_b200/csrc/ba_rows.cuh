// ba_rows.cuh — K4/K5 with a two-kernel, atomic-free, bit-reproducible camera-side reduction (scatter mode 4, the default).
//
//   K5a  k_schur_rows   : one small CTA per tile (same shape as K7, which streams J at ~75 % of the HBM roofline): loads J once,
//                         per-point reduce through shared memory, then the tile's 6-vectors are summed per camera inside the
//                         CTA (static camera-sorted schedule, two-level serial sum) and each (tile, camera) total is written as
//                         one 64-byte row of a compact partial buffer.  No atomics, no read-modify-write, no persistent state.
//   K5b  k_camera_gather: one warp per camera adds that camera's rows (static list), applies Df^2 p and forms p.q for PCG.
//
// Row traffic is 64 B per distinct (tile, camera) pair each way: with the locality real SfM maps have (points numbered in
// the order their first camera was added) that is a fraction of the observation count.  
#pragma once
#include "ba_scatter.cuh"

namespace bak {

constexpr int ROW_DOUBLES = 8;    // 6 used; 64-byte rows = two full 32-byte sectors
constexpr int L1_CHUNK = 8;       // level-1 chunk of the in-tile segmented sum

struct RowSched {
    const unsigned char* slot;    // [No] rank of the observation in its tile's camera-sorted order
    const unsigned char* lvl1;    // [No] per sorted position: rows this thread adds up in level 1 (0: none)
    const unsigned char* hrank;   // [No] per sorted position: index of this segment head among the tile's heads
    const int* spack;             // [No] per sorted position: camera | (segment length << 22) at segment heads, -1 elsewhere
    const int* tile_hbase;        // [nTiles+1] first compact row of each tile
    double* rows;                 // [nRows][ROW_DOUBLES]
    double* tile_intr;            // [nTiles][4] per-tile intrinsics-side partials (free intrinsics, one group)
    // by-camera view of the rows
    const int* cam_row_off;       // [Nc+1]
    const int* cam_rows;          // [nRows] compact row ids, ascending
    int nRows;
};

template <int MODE, bool FI>
__global__ void __launch_bounds__(TILE_OBS)
k_schur_rows(View V, RowSched S, const double* __restrict__ xin, double* __restrict__ atomic_out, const int* __restrict__ done_flag) {
    __shared__ double sW[TILE_OBS][3];
    __shared__ double sV[TILE_OBS][3];
    __shared__ __align__(16) double sC[TILE_OBS][6];
    __shared__ double sred[32];
    if (done_flag && *done_flag) return;
    const int t = blockIdx.x;
    const int o0 = V.tile_off[t], o1 = V.tile_off[t + 1];
    const size_t ld = V.ldo;
    const int ioff = 6 * V.Nc;
    double yi[4] = {0, 0, 0, 0};

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
                double W0 = 0, W1 = 0, W2 = 0;
                for (int i = V.pt_off[pj] - o0, e = V.pt_off[pj + 1] - o0; i < e; ++i) { W0 += sW[i][0]; W1 += sW[i][1]; W2 += sW[i][2]; }
                const double* H = V.Hinv + 6 * size_t(pj);
                sV[threadIdx.x][0] = H[0] * W0 + H[1] * W1 + H[2] * W2;
                sV[threadIdx.x][1] = H[1] * W0 + H[3] * W1 + H[4] * W2;
                sV[threadIdx.x][2] = H[2] * W0 + H[4] * W1 + H[5] * W2;
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
        int n1 = 0, pk = -1;
        if (act) {
            double2* sc = reinterpret_cast<double2*>(sC[S.slot[o]]);
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
            n1 = S.lvl1[o];
            pk = S.spack[o];
        }
        __syncthreads();
        // level 1: chunk leaders (segment heads and every 8th sorted position) fold their <= 8 rows into their own row
        if (n1 > 1) {
            double2* row = reinterpret_cast<double2*>(sC[threadIdx.x]);
            double2 a = row[0], bq = row[1], cq = row[2];
            for (int i = 1; i < n1; ++i) {
                const double2 a2 = row[3 * i], b2 = row[3 * i + 1], c2 = row[3 * i + 2];
                a.x += a2.x; a.y += a2.y; bq.x += b2.x; bq.y += b2.y; cq.x += c2.x; cq.y += c2.y;
            }
            row[0] = a; row[1] = bq; row[2] = cq;
        }
        __syncthreads();
        // level 2: each segment head adds the chunk leaders of its segment and writes one compact 64-byte row
        if (pk >= 0) {
            const int len = pk >> 22;
            const int j = threadIdx.x, end = j + len;
            const double2* row = reinterpret_cast<const double2*>(sC[j]);
            double2 a = row[0], bq = row[1], cq = row[2];
            for (int q = (j / L1_CHUNK + 1) * L1_CHUNK; q < end; q += L1_CHUNK) {
                const double2* r2 = reinterpret_cast<const double2*>(sC[q]);
                const double2 a2 = r2[0], b2 = r2[1], c2 = r2[2];
                a.x += a2.x; a.y += a2.y; bq.x += b2.x; bq.y += b2.y; cq.x += c2.x; cq.y += c2.y;
            }
            double2* dst = reinterpret_cast<double2*>(S.rows + size_t(ROW_DOUBLES) * (S.tile_hbase[t] + S.hrank[o]));
            __stcg(dst, a); __stcg(dst + 1, bq); __stcg(dst + 2, cq);
        }
    } else {
        // ---- one point with more than TILE_OBS observations (rare): two passes, every observation is its own row ----
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
        const int hb = S.tile_hbase[t];
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
            double2* dst = reinterpret_cast<double2*>(S.rows + size_t(ROW_DOUBLES) * (hb + (o - o0)));
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
    if (FI && V.Ng == 1) {   // one shared intrinsics group: per-tile partial, summed in tile order by k_camera_gather
#pragma unroll
        for (int k = 0; k < 4; ++k) { const double s = block_sum(yi[k], sred); if (threadIdx.x == 0) S.tile_intr[4 * size_t(t) + k] = s; }
    }
}

// K5b — y[c] = sum of camera c's partial rows (one warp per camera, lanes stride the static row list, fixed shuffle tree);
// with do_dot: q = y + Df2 p stored in y, block partials of p.q, last block forms alpha = rho / p.q  (PcgState2).
// Block `gridDim.x - 1` additionally reduces the per-tile intrinsics partials (free intrinsics, single group).
constexpr int GATHER_WARPS = 4;
__global__ void __launch_bounds__(GATHER_WARPS * 32)
k_camera_gather(int Nc, int Ng, int fi, int nTiles, RowSched S, const double* __restrict__ Df2, const double* __restrict__ p,
                double* __restrict__ y, double* __restrict__ partial, PcgState2* __restrict__ st, int do_dot) {
    __shared__ double sw[GATHER_WARPS];
    __shared__ double sred[32];
    __shared__ bool last;
    if (do_dot && st->done) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int ioff = 6 * Nc;
    double dot = 0.0;
    const bool intr_block = (blockIdx.x == gridDim.x - 1);
    if (!intr_block) {
        const int c = blockIdx.x * GATHER_WARPS + wid;
        if (c < Nc) {
            double s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0, s5 = 0;
            for (int i = S.cam_row_off[c] + lane, e = S.cam_row_off[c + 1]; i < e; i += 32) {
                const double2* row = reinterpret_cast<const double2*>(S.rows + size_t(ROW_DOUBLES) * S.cam_rows[i]);
                const double2 a = __ldcg(row), bq = __ldcg(row + 1), cq = __ldcg(row + 2);
                s0 += a.x; s1 += a.y; s2 += bq.x; s3 += bq.y; s4 += cq.x; s5 += cq.y;
            }
            s0 = warp_sum(s0); s1 = warp_sum(s1); s2 = warp_sum(s2); s3 = warp_sum(s3); s4 = warp_sum(s4); s5 = warp_sum(s5);
            if (lane < 6) {
                double v = lane == 0 ? s0 : lane == 1 ? s1 : lane == 2 ? s2 : lane == 3 ? s3 : lane == 4 ? s4 : s5;
                const int j = 6 * c + lane;
                if (do_dot) { const double pj = p[j]; v += Df2[j] * pj; dot = pj * v; }
                y[j] = v;
            }
        }
    } else if (fi) {
        // intrinsics entries: single group -> ordered sum of the per-tile partials; several groups -> RED-accumulated in y already
        for (int k = wid; k < 4 * Ng; k += GATHER_WARPS) {
            double v;
            if (Ng == 1) {
                double s = 0.0;
                for (int t = lane; t < nTiles; t += 32) s += __ldcg(S.tile_intr + 4 * size_t(t) + k);
                v = warp_sum(s);
            } else v = y[ioff + k];
            if (lane == 0) {
                const int j = ioff + k;
                if (do_dot) { const double pj = p[j]; v += Df2[j] * pj; dot += pj * v; }
                y[j] = v;
            }
        }
    }
    if (!do_dot) return;
    dot = warp_sum(dot);
    if (lane == 0) sw[wid] = dot;
    __syncthreads();
    if (threadIdx.x == 0) {
        double bsum = 0.0;
        for (int w = 0; w < GATHER_WARPS; ++w) bsum += sw[w];
        partial[blockIdx.x] = bsum;
        __threadfence();
        const unsigned int tk = atomicAdd(&st->ticket_q, 1u);
        last = (tk == gridDim.x - 1);
    }
    __syncthreads();
    if (last) {   // fixed-order sum of the block partials by one warp
        __threadfence();
        double pq = 0.0;
        if (wid == 0) {
            for (int i = lane; i < int(gridDim.x); i += 32) pq += __ldcg(partial + i);
            pq = warp_sum(pq);
            if (lane == 0) {
                if (!(pq > 0.0) || !isfinite(pq)) { st->done = 3; st->alpha = 0.0; }
                else st->alpha = st->rho / pq;
                st->ticket_q = 0;
            }
        }
    }
    (void)sred;
}

}  // namespace bak
