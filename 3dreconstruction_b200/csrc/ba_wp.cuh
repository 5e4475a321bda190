// ba_wp.cuh — K4 / K5a on WARP-PACKED tiles (scatter_mode 7).
//
// Same partial-row formulation as ba_rows.cuh (per-(tile, camera) rows + by-camera gather, no atomics, bit-reproducible), with two changes
// that ncu asked for on k_schur_rows (0.72 of the measured HBM peak against 0.90 for K4, the same streams without the per-point phase):
//   * the per-point phase never leaves the warp.  A tile is 8 warps x 32 lanes and every warp holds WHOLE points (a point's observations
//     are contiguous lanes), so  W_p = sum Jp^T u  is a warp-segmented shuffle reduction to the point's first lane, that lane applies
//     Hinv_p, and  v = Hinv_p W_p  goes back to the point's lanes with one indexed shuffle: no shared memory, no CTA barrier, and the warps
//     of a CTA run independently until the one barrier of the camera fold (k_schur_rows has three barriers and two shared-memory round
//     trips per tile);
//   * 256 positions per tile instead of 128: a camera's observations inside a tile fold into fewer partial rows (rows are what K5a writes
//     and the gather reads back every PCG iteration).
// A point with more than 32 observations cannot live in one warp: it becomes a tile of its own and takes the generic two-pass path
// (tile_rows_long).  The schedule (ba_engine.cu: build_wp_schedule) is static: 12 bytes per position.
#pragma once
#include "ba_rows.cuh"

namespace bak {

#ifndef BA_WP_THREADS
#define BA_WP_THREADS 256             // positions (threads) per tile: 256 or 128 (A/B knob)
#endif
constexpr int WP_THREADS = BA_WP_THREADS;
constexpr int WP_WARPS = WP_THREADS / 32;
constexpr int WP_LONG = 1 << 30;      // tile_info2.x flag: the tile is one point with more than 32 observations

#ifndef BA_WP_MINB
#define BA_WP_MINB (768 / BA_WP_THREADS)   // CTAs/SM for ptxas: 768 threads = 80 registers
#endif

template <int MODE, bool FI, int PK>
__global__ void __launch_bounds__(WP_THREADS, BA_WP_MINB)
k_schur_wp(View V, RowSched S, const double* __restrict__ xin, double* __restrict__ atomic_out, const int* __restrict__ done_flag) {
    __shared__ __align__(16) double sC[WP_THREADS][6];
    __shared__ double sred[8];
    pdl_trigger();
    const int t = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int4 ti = __ldg(S.tile_info2 + t);
    const int hbase = ti.y;
    const int ioff = 6 * V.Nc;
    double yi[4] = {0, 0, 0, 0};
    if (ti.x & WP_LONG) {
        pdl_wait();
        if (done_flag && *done_flag) return;
        tile_rows_long<MODE, FI, PK, WP_THREADS>(V, S, sred, BarCta{}, tid, ti.z, ti.x & (WP_LONG - 1), hbase, ti.w, xin, atomic_out, yi);
    } else {
        // ---- every global load of this tile that does not depend on the predecessor kernel is issued here ----
        const int n_tile = ti.x;
        const int4 wi = __ldg(S.winfo + size_t(t) * WP_WARPS + warp);
        const bool act = lane < (wi.y & 0xff);
        const int o = wi.x + lane;
        const int2 ra = __ldg(S.wpA + size_t(t) * WP_THREADS + tid);
        const int pk = __ldg(S.wpB + size_t(t) * WP_THREADS + tid);
        const int c = ra.x & 0x3fffff, slot = int(unsigned(ra.x) >> 24);
        const int head = ra.y & 31, segrem = (ra.y >> 5) & 31, prank = (ra.y >> 10) & 31, rem = (ra.y >> 15) & 31, hrank = (ra.y >> 20) & 0xff;
        double2 jc[6], jp[3], ji[4], rr = make_double2(0.0, 0.0);
#pragma unroll
        for (int k = 0; k < 6; ++k) jc[k] = make_double2(0.0, 0.0);
#pragma unroll
        for (int k = 0; k < 3; ++k) jp[k] = make_double2(0.0, 0.0);
        if (act) {
            const double2* J = V.J2 + o;
            load_jc<PK>(V, J, jc);
            load_jp<PK>(V, J, jp);
            if (FI) load_ji<PK>(V, J, ji);
            if (MODE == SWEEP_RHS) rr = V.r2[o];
        }
        const bool is_head = act && head == lane;
        double2 h0 = make_double2(0.0, 0.0), h1 = h0, h2 = h0;
        if (MODE == SWEEP_MATVEC && is_head) {
            const double2* H = reinterpret_cast<const double2*>(V.Hinv + 6 * size_t(wi.z + prank));
            h0 = __ldg(H); h1 = __ldg(H + 1); h2 = __ldg(H + 2);      // (H0,H1) (H2,H3) (H4,H5)
        }
        // everything above is static during a linear solve; xin and the done flag are the predecessor's (PCG tail) output
        pdl_wait();
        if (done_flag && *done_flag) return;
        const int g = (FI && act) ? V.cam_group[c] : 0;
        double2 u = make_double2(0.0, 0.0), tt = make_double2(0.0, 0.0);
        if (MODE == SWEEP_MATVEC) {
            double W0 = 0.0, W1 = 0.0, W2 = 0.0;
            if (act) {
                const double2* xc = reinterpret_cast<const double2*>(xin + 6 * c);
                const double2 x0 = __ldg(xc), x1 = __ldg(xc + 1), x2 = __ldg(xc + 2);
                u.x = jc[0].x * x0.x + jc[1].x * x0.y + jc[2].x * x1.x + jc[3].x * x1.y + jc[4].x * x2.x + jc[5].x * x2.y;
                u.y = jc[0].y * x0.x + jc[1].y * x0.y + jc[2].y * x1.x + jc[3].y * x1.y + jc[4].y * x2.x + jc[5].y * x2.y;
                if (FI) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) { const double xv = __ldg(xin + ioff + 4 * g + k); u.x += ji[k].x * xv; u.y += ji[k].y * xv; }
                }
                W0 = jp[0].x * u.x + jp[0].y * u.y; W1 = jp[1].x * u.x + jp[1].y * u.y; W2 = jp[2].x * u.x + jp[2].y * u.y;
            }
            // segmented reduction to the first lane of every point: after the step with distance d a lane holds the sum over the lanes
            // [lane, min(lane + 2d - 1, last lane of its point)]; only as many steps as the longest point of this warp needs
            const int maxseg = __reduce_max_sync(0xffffffffu, act ? segrem : 0);
            for (int d = 1; d <= maxseg; d <<= 1) {
                const double a0 = __shfl_down_sync(0xffffffffu, W0, d), a1 = __shfl_down_sync(0xffffffffu, W1, d), a2 = __shfl_down_sync(0xffffffffu, W2, d);
                if (d <= segrem) { W0 += a0; W1 += a1; W2 += a2; }
            }
            double v0 = h0.x * W0 + h0.y * W1 + h1.x * W2;      // Hinv rows (H0,H1,H2) (H1,H3,H4) (H2,H4,H5); meaningful in head lanes
            double v1 = h0.y * W0 + h1.y * W1 + h2.x * W2;
            double v2 = h1.x * W0 + h2.x * W1 + h2.y * W2;
            v0 = __shfl_sync(0xffffffffu, v0, head); v1 = __shfl_sync(0xffffffffu, v1, head); v2 = __shfl_sync(0xffffffffu, v2, head);
            tt.x = u.x - (jp[0].x * v0 + jp[1].x * v1 + jp[2].x * v2);
            tt.y = u.y - (jp[0].y * v0 + jp[1].y * v1 + jp[2].y * v2);
        } else if (act) {   // SWEEP_RHS: t = r - E vp
            const double* v = V.vp + 3 * size_t(wi.z + prank);
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
        // camera fold: sorted position tid; one warp per 32 sorted positions, segmented shuffle reduction over the runs of one camera
        double2 a = make_double2(0.0, 0.0), bq = a, cq = a;
        if (tid < n_tile) {
            const double2* row = reinterpret_cast<const double2*>(sC[tid]);
            a = row[0]; bq = row[1]; cq = row[2];
        }
        const int maxrem = __reduce_max_sync(0xffffffffu, rem);
        for (int d = 1; d <= maxrem; d <<= 1) {
            const double ax = __shfl_down_sync(0xffffffffu, a.x, d), ay = __shfl_down_sync(0xffffffffu, a.y, d);
            const double bx = __shfl_down_sync(0xffffffffu, bq.x, d), by = __shfl_down_sync(0xffffffffu, bq.y, d);
            const double cx = __shfl_down_sync(0xffffffffu, cq.x, d), cy = __shfl_down_sync(0xffffffffu, cq.y, d);
            if (d <= rem) { a.x += ax; a.y += ay; bq.x += bx; bq.y += by; cq.x += cx; cq.y += cy; }
        }
        if (pk >= 0) {
            double2* dst = reinterpret_cast<double2*>(S.rows + size_t(ROW_DOUBLES) * (hbase + hrank));
            row_store(dst, a); row_store(dst + 1, bq); row_store(dst + 2, cq);
        }
    }
    if (FI && V.Ng == 1) tile_intr_store<WP_THREADS>(S, sred, BarCta{}, tid, t, yi);
}

}  // namespace bak
