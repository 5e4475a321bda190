// ba_rows.cuh — K4 (reduced right-hand side) and K5 (implicit Schur matvec, the PCG inner loop) with an atomic-free, bit-reproducible
// camera-side reduction: per-(tile, camera) partial rows + a by-camera gather.  
//
//   K5a, two launch shapes over the SAME tile math (tile_rows below), the same static schedule and bit-identical rows:
//     k_schur_rows   (scatter_mode 4): one CTA of 128 threads per tile; all global loads of the tile issued at kernel entry.
//     k_schur_stream (scatter_mode 6, ba_stream.cuh): persistent, warp-specialised, tiles streamed through a shared-memory ring with
//                    cp.async.bulk + mbarrier.
//   K5b  k_camera_gather: one CTA per camera adds that camera's rows in the fixed order of a static list.
//   Pose-factored forms (the defaults, second half of this file): k_pose_factors (G_c per camera), k_schur_rows_f (K5a and
//   K4 from the point columns only), k_backsub_f + k_pose_apply (K7), k2c_camera_blocks_f (K2c); the gather multiplies by G_c^T.
//
// Tile math: thread i = observation i of the tile.  u = Jc p[c] (+ Ji p[g]);  per point W = sum Jp^T u through shared memory (points
// are contiguous, thread j sums point j), v = Hinv W;  t = u - Jp v;  the 6-vector Jc^T t is written to shared memory at the
// observation's rank in the tile's static camera-sorted order; each warp then folds 32 sorted positions with a segmented shuffle
// reduction (fixed tree) and the head of every run of one camera stores one packed 48-byte row.  No atomics, no read-modify-write.
#pragma once
#include "ba_pcg.cuh"

namespace bak {

#ifndef BA_ROW_DOUBLES
#define BA_ROW_DOUBLES 6
#endif
constexpr int ROW_DOUBLES = BA_ROW_DOUBLES;    // 6: packed 48-byte rows; 8: 64-byte rows (two whole sectors; measured 5 % slower end to end: the pad is written every iteration)
constexpr int GATHER_WARPS = 4;                // warps per CTA of the by-camera gather (K5b)
constexpr int TILE_MAX_PTS = 64;               // points per tile (the streaming kernel reserves 48 bytes of Hinv per point in every stage)

// L2 residency (BA_L2_HINTS): the partial rows of one PCG iteration (108 MB on the Venice shape) are written by K5a and read back by the
// gather a few microseconds later, and overwritten by the next K5a; in between K5a streams 3-8x as many read-once bytes through the same
// L2 (126 MB), which evicts the rows before the gather gets to them (ncu: 37 % L2 hit rate in the gather).  With a cache policy the rows
// are kept (evict_last) and the read-once streams of the pose-factored K5a go first (evict_first).
#ifndef BA_L2_HINTS
#define BA_L2_HINTS 0          // bit 0: read-once streams evict_first, bit 1: row stores evict_last, bit 2: the gather's row loads evict_last
#endif
__device__ __forceinline__ uint64_t l2_policy_evict_first() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ uint64_t l2_policy_evict_last() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ void row_store(double2* p, double2 v) { __stcg(p, v); }
__device__ __forceinline__ double2 row_load(const double2* p) { return __ldcg(p); }
__device__ __forceinline__ void row_store_keep(double2* p, double2 v, uint64_t pol) {
#if BA_L2_HINTS & 2
    asm volatile("st.global.cg.L2::cache_hint.v2.f64 [%0], {%1, %2}, %3;" ::"l"(p), "d"(v.x), "d"(v.y), "l"(pol) : "memory");
#else
    __stcg(p, v);
#endif
}
__device__ __forceinline__ double2 row_load_keep(const double2* p, uint64_t pol) {
#if BA_L2_HINTS & 4
    double2 v;
    asm volatile("ld.global.cg.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;" : "=d"(v.x), "=d"(v.y) : "l"(p), "l"(pol) : "memory");
    return v;
#else
    return __ldcg(p);
#endif
}
__device__ __forceinline__ double2 ldg2_once(const double2* p, uint64_t pol) {      // read-once stream
#if BA_L2_HINTS & 1
    double2 v;
    asm volatile("ld.global.nc.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;" : "=d"(v.x), "=d"(v.y) : "l"(p), "l"(pol));
    return v;
#else
    return __ldg(p);
#endif
}
__device__ __forceinline__ int4 ldg4_once(const int4* p, uint64_t pol) {
#if BA_L2_HINTS & 1
    int4 v;
    asm volatile("ld.global.nc.L2::cache_hint.v4.s32 {%0, %1, %2, %3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol));
    return v;
#else
    return __ldg(p);
#endif
}

struct RowSched {
    // One 16-byte record per tile position i = o0 + tid:
    //   .x  camera of observation i | sorted position of the head of its camera segment << 24
    //   .y  slot | rem << 8 | hrank << 16 | lpt << 24   (observation i: rank in the tile's camera-sorted order; sorted position
    //       tid: positions of its run after it, row number of the run in the tile; observation i: tile-local point)
    //   .z  sorted position tid: camera | run length << 22 at the head of a run, -1 elsewhere  (run = maximal stretch of one camera
    //       inside a warp's 32 sorted positions; one partial row per run)
    //   .w  tile-local point tid: first observation | observation count << 16 (0 when tid >= points in tile)
    const int4* tmeta;            // [No]
    const int4* tile_info;        // [nTiles] {o0, n | points << 24 (points = 0 for a long-point tile), first compact row, first local point}
    double* rows;                 // [nRows][ROW_DOUBLES]
    double* tile_intr;            // [nTiles][4] per-tile intrinsics-side partials (free intrinsics, one group)
    const int* cam_row_off;       // [Nc+1] by-camera view of the rows
    const int* cam_rows;          // [nRows] compact row ids, ascending
    int nRows;
};

// barrier policies of the tile math: the whole CTA (k_schur_rows) or one 128-thread consumer group (k_schur_stream)
struct BarCta { __device__ __forceinline__ void operator()() const { __syncthreads(); } };
struct BarGroup {
    int id;
    __device__ __forceinline__ void operator()() const { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(TILE_OBS) : "memory"); }
};

struct TileScratch {   // 12 KB per tile in flight
    double sW[TILE_OBS][3];
    double sV[TILE_OBS][3];
    alignas(16) double sC[TILE_OBS][6];
    double sred[8];
};

// Sum over the NT threads of a tile (fixed tree); valid in every thread.
template <int NT = TILE_OBS, class Bar>
__device__ __forceinline__ double tile_sum(double v, double* sred, Bar bar, int tid) {
    v = warp_sum(v);
    bar();
    if ((tid & 31) == 0) sred[tid >> 5] = v;
    bar();
    if (NT == 128) return (sred[0] + sred[1]) + (sred[2] + sred[3]);
    return ((sred[0] + sred[1]) + (sred[2] + sred[3])) + ((sred[4] + sred[5]) + (sred[6] + sred[7]));
}

// The operands of one tile position, in registers.
template <bool FI>
struct TileRegs {
    int4 tm;
    double2 jc[6], jp[3], ji[4], rr;   // ji: free intrinsics only (dead otherwise)
    double2 h0, h1, h2;      // Hinv of tile-local point tid (MATVEC)
};

// Tile math from registers to the tile's partial rows.  tid = position in the tile (0..127), n = observations of the tile.
template <int MODE, bool FI, class Bar>
__device__ __forceinline__ void tile_rows(const View& V, const RowSched& S, TileScratch& M, Bar bar, const int tid, const int n, const int hbase,
                                          const int pfirst, const TileRegs<FI>& R, const double* __restrict__ xin, double* __restrict__ atomic_out,
                                          double (&yi)[4]) {
    const bool act = tid < n;
    const int4 tm = R.tm;
    const int npt_obs = tm.w >> 16, pstart = tm.w & 0xffff;   // this thread as a point thread
    const int c = tm.x & 0xffffff;
    const int slot = tm.y & 0xff, n1 = (tm.y >> 8) & 0xff, hrank = (tm.y >> 16) & 0xff, lpt = (tm.y >> 24) & 0xff;
    const int pk = tm.z;
    const int ioff = 6 * V.Nc;
    const int g = (FI && act) ? V.cam_group[c] : 0;
    double2 u = make_double2(0.0, 0.0), tt = make_double2(0.0, 0.0);
    if (MODE == SWEEP_MATVEC) {
        if (act) {
#ifdef BA_EXP_NO_PGATHER      // timing experiment only (wrong results): what the divergent gather of p[c] costs
            const double2 x0 = make_double2(1.0, 0.5), x1 = make_double2(0.25, 2.0), x2 = make_double2(1.5, 0.75);
#else
            const double2* xc = reinterpret_cast<const double2*>(xin + 6 * c);
            const double2 x0 = __ldg(xc), x1 = __ldg(xc + 1), x2 = __ldg(xc + 2);      // (one 32-byte + one 16-byte load by the parity of c measured 1 % slower)
#endif
            u.x = R.jc[0].x * x0.x + R.jc[1].x * x0.y + R.jc[2].x * x1.x + R.jc[3].x * x1.y + R.jc[4].x * x2.x + R.jc[5].x * x2.y;
            u.y = R.jc[0].y * x0.x + R.jc[1].y * x0.y + R.jc[2].y * x1.x + R.jc[3].y * x1.y + R.jc[4].y * x2.x + R.jc[5].y * x2.y;
            if (FI) {
#pragma unroll
                for (int k = 0; k < 4; ++k) { const double xv = __ldg(xin + ioff + 4 * g + k); u.x += R.ji[k].x * xv; u.y += R.ji[k].y * xv; }
            }
        }
#ifndef BA_EXP_NO_POINT      // timing experiment only (wrong results): what the per-point phase (two barriers, shared-memory exchange) costs
        // W_p = sum over the point's observations of Jp^T u.  First inside the warp: the lanes of one point are contiguous, so a segmented
        // shuffle reduction (fixed tree, only as many steps as the longest run in this warp needs) leaves the sum of every run in its first
        // lane; only those lanes go through shared memory, and the point thread adds one value per warp its point touches (1.1 on average)
        // instead of walking all its observations serially (5 on average, up to 40: the tile's critical path between the two barriers).
        {
            double w0 = 0.0, w1 = 0.0, w2 = 0.0;
            if (act) { w0 = R.jp[0].x * u.x + R.jp[0].y * u.y; w1 = R.jp[1].x * u.x + R.jp[1].y * u.y; w2 = R.jp[2].x * u.x + R.jp[2].y * u.y; }
            const unsigned same = __match_any_sync(0xffffffffu, act ? lpt : 0xffff);      // lanes of my point in this warp (contiguous)
            const int lane = tid & 31;
            const int segrem = __popc(same >> lane) - 1;                                   // lanes of the run after this one
            const int maxseg = __reduce_max_sync(0xffffffffu, segrem);
            for (int d = 1; d <= maxseg; d <<= 1) {
                const double a0 = __shfl_down_sync(0xffffffffu, w0, d), a1 = __shfl_down_sync(0xffffffffu, w1, d), a2 = __shfl_down_sync(0xffffffffu, w2, d);
                if (d <= segrem) { w0 += a0; w1 += a1; w2 += a2; }
            }
            if (act && (same & ((1u << lane) - 1u)) == 0) { M.sW[tid][0] = w0; M.sW[tid][1] = w1; M.sW[tid][2] = w2; }   // first lane of its run
        }
        bar();
        if (npt_obs > 0) {
            double W0 = M.sW[pstart][0], W1 = M.sW[pstart][1], W2 = M.sW[pstart][2];
            for (int i = (pstart / 32 + 1) * 32, e = pstart + npt_obs; i < e; i += 32) { W0 += M.sW[i][0]; W1 += M.sW[i][1]; W2 += M.sW[i][2]; }
            M.sV[tid][0] = R.h0.x * W0 + R.h0.y * W1 + R.h1.x * W2;      // Hinv rows (H0,H1,H2) (H1,H3,H4) (H2,H4,H5)
            M.sV[tid][1] = R.h0.y * W0 + R.h1.y * W1 + R.h2.x * W2;
            M.sV[tid][2] = R.h1.x * W0 + R.h2.x * W1 + R.h2.y * W2;
        }
        bar();
        if (act) {
            const double* v = M.sV[lpt];
            tt.x = u.x - (R.jp[0].x * v[0] + R.jp[1].x * v[1] + R.jp[2].x * v[2]);
            tt.y = u.y - (R.jp[0].y * v[0] + R.jp[1].y * v[1] + R.jp[2].y * v[2]);
        }
#else
        if (act) { tt.x = u.x - (R.jp[0].x * R.h0.x + R.jp[1].x * R.h1.x + R.jp[2].x * R.h2.x); tt.y = u.y - (R.jp[0].y * R.h0.y + R.jp[1].y * R.h1.y + R.jp[2].y * R.h2.y); }
#endif
    } else if (act) {   // SWEEP_RHS: t = r - E vp
        const double* v = V.vp + 3 * size_t(pfirst + lpt);
        tt.x = R.rr.x - (R.jp[0].x * v[0] + R.jp[1].x * v[1] + R.jp[2].x * v[2]);
        tt.y = R.rr.y - (R.jp[0].y * v[0] + R.jp[1].y * v[1] + R.jp[2].y * v[2]);
    }
    if (act) {
        double2* sc = reinterpret_cast<double2*>(M.sC[slot]);
        sc[0] = make_double2(R.jc[0].x * tt.x + R.jc[0].y * tt.y, R.jc[1].x * tt.x + R.jc[1].y * tt.y);
        sc[1] = make_double2(R.jc[2].x * tt.x + R.jc[2].y * tt.y, R.jc[3].x * tt.x + R.jc[3].y * tt.y);
        sc[2] = make_double2(R.jc[4].x * tt.x + R.jc[4].y * tt.y, R.jc[5].x * tt.x + R.jc[5].y * tt.y);
        if (FI) {
#pragma unroll
            for (int k = 0; k < 4; ++k) yi[k] = R.ji[k].x * tt.x + R.ji[k].y * tt.y;
            if (V.Ng > 1) {
#pragma unroll
                for (int k = 0; k < 4; ++k) red_add(atomic_out + ioff + 4 * g + k, yi[k]);
            }
        }
    }
    bar();
    // Fold over the camera-sorted rows, one warp per 32 sorted positions: a segmented shuffle reduction (fixed tree => bit-reproducible).
    // After the step with distance d a lane holds the sum of the rows [lane, min(lane + 2d - 1, end of its run)]; `rem` = positions of
    // the run after this one (static schedule).  The head of every run stores one compact row.  (The first version folded serially
    // through shared memory in two levels: ncu showed 60 % of the kernel's shared-memory wavefronts and its longest dependent chains there.)
    {
        const int rem = n1;
        double2 a = make_double2(0.0, 0.0), bq = a, cq = a;
        if (act) {
            const double2* row = reinterpret_cast<const double2*>(M.sC[tid]);
            a = row[0]; bq = row[1]; cq = row[2];
        }
        const int maxrem = __reduce_max_sync(0xffffffffu, act ? rem : 0);   // shuffles share the shared-memory data pipe: only the steps the longest run needs
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
}

// One point with more than TILE_OBS observations (rare): two passes straight from global memory, every observation is its own row.
// Whole tile group (128 threads) calls; p = the point.  (Inlined on purpose: as a real call it takes the View by reference, which makes
// every thread of every CTA spill the whole kernel-parameter struct to local memory at kernel entry — 25 STL per thread, measured
// +30 % on K4/K5a.)
template <int MODE, bool FI, int PK, int NT = TILE_OBS, class Bar>
__device__ __forceinline__ void tile_rows_long(const View& V, const RowSched& S, double* sred, Bar bar, const int tid, const int o0, const int n,
                                            const int hbase, const int p, const double* __restrict__ xin, double* __restrict__ atomic_out,
                                            double (&yi)[4]) {
    const int ioff = 6 * V.Nc;
    const int o1 = o0 + n;
    double v[3];
    if (MODE == SWEEP_MATVEC) {
        double W[3] = {0, 0, 0};
        for (int o = o0 + tid; o < o1; o += NT) {
            const int c = V.obs_cam[o];
            const double2* J = V.J2 + o;
            double2 u = make_double2(0.0, 0.0);
            double2 jc[6], jp[3];
            load_jc<PK>(V, J, jc);
#pragma unroll
            for (int k = 0; k < 6; ++k) { const double xv = __ldg(xin + 6 * c + k); u.x += jc[k].x * xv; u.y += jc[k].y * xv; }
            if (FI) {
                const int g = V.cam_group[c];
                double2 ji[4];
                load_ji<PK>(V, J, ji);
#pragma unroll
                for (int k = 0; k < 4; ++k) { const double xv = __ldg(xin + ioff + 4 * g + k); u.x += ji[k].x * xv; u.y += ji[k].y * xv; }
            }
            load_jp<PK>(V, J, jp);
#pragma unroll
            for (int a = 0; a < 3; ++a) W[a] += jp[a].x * u.x + jp[a].y * u.y;
        }
        double w[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) w[a] = tile_sum<NT>(W[a], sred, bar, tid);
        const double* H = V.Hinv + 6 * size_t(p);
        v[0] = H[0] * w[0] + H[1] * w[1] + H[2] * w[2]; v[1] = H[1] * w[0] + H[3] * w[1] + H[4] * w[2]; v[2] = H[2] * w[0] + H[4] * w[1] + H[5] * w[2];
    } else { v[0] = V.vp[3 * size_t(p)]; v[1] = V.vp[3 * size_t(p) + 1]; v[2] = V.vp[3 * size_t(p) + 2]; }
    for (int o = o0 + tid; o < o1; o += NT) {
        const int c = V.obs_cam[o];
        const int g = FI ? V.cam_group[c] : 0;
        const double2* J = V.J2 + o;
        double2 jc[6], ji[4], jp[3];
        double2 u = make_double2(0.0, 0.0);
        load_jc<PK>(V, J, jc);
        if (FI) load_ji<PK>(V, J, ji);
        if (MODE == SWEEP_MATVEC) {
#pragma unroll
            for (int k = 0; k < 6; ++k) { const double xv = __ldg(xin + 6 * c + k); u.x += jc[k].x * xv; u.y += jc[k].y * xv; }
            if (FI) {
#pragma unroll
                for (int k = 0; k < 4; ++k) { const double xv = __ldg(xin + ioff + 4 * g + k); u.x += ji[k].x * xv; u.y += ji[k].y * xv; }
            }
        }
        double e0 = 0.0, e1 = 0.0;
        load_jp<PK>(V, J, jp);
#pragma unroll
        for (int a = 0; a < 3; ++a) { e0 += jp[a].x * v[a]; e1 += jp[a].y * v[a]; }
        double2 tt;
        if (MODE == SWEEP_MATVEC) { tt.x = u.x - e0; tt.y = u.y - e1; }
        else { const double2 r = V.r2[o]; tt.x = r.x - e0; tt.y = r.y - e1; }
        double2* dst = reinterpret_cast<double2*>(S.rows + size_t(ROW_DOUBLES) * (hbase + (o - o0)));
        row_store(dst, make_double2(jc[0].x * tt.x + jc[0].y * tt.y, jc[1].x * tt.x + jc[1].y * tt.y));
        row_store(dst + 1, make_double2(jc[2].x * tt.x + jc[2].y * tt.y, jc[3].x * tt.x + jc[3].y * tt.y));
        row_store(dst + 2, make_double2(jc[4].x * tt.x + jc[4].y * tt.y, jc[5].x * tt.x + jc[5].y * tt.y));
        if (FI) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const double contrib = ji[k].x * tt.x + ji[k].y * tt.y;
                if (V.Ng > 1) red_add(atomic_out + ioff + 4 * g + k, contrib); else yi[k] += contrib;
            }
        }
    }
}

// One shared intrinsics group: the tile's partial of the intrinsics entries, summed in tile order by k_camera_gather.
template <int NT = TILE_OBS, class Bar>
__device__ __forceinline__ void tile_intr_store(const RowSched& S, double* sred, Bar bar, int tid, int t, const double (&yi)[4]) {
#pragma unroll
    for (int k = 0; k < 4; ++k) { const double s = tile_sum<NT>(yi[k], sred, bar, tid); if (tid == 0) S.tile_intr[4 * size_t(t) + k] = s; }
}

// ---------------------------------------------------------------------------------------------------------
// Pose-factored Schur product.  Both models reach the pose only through P = R X + t, so with A = d(residual)/dP
// the stored point columns are Jp = w A R and the pose columns are Jc = w A dP/d(pose) = Jp K(X) G_c with
//     K(X) = [ -[X]x | I ]   (3x6, the POINT in world coordinates),   K(X) (v; tau) = v x X + tau,
//     G_c  = [ J_r(aa) 0 ; 0 R^T ]                      angle-axis + translation (Ceres policy; J_r = right Jacobian of SO(3)),
//     G_c  = [ R^T 0 ; -[R^T t]x R^T  R^T ]             left twist (g2o policy),        G_c = 0 for a constant pose.
// The PCG matvec therefore needs, per observation, only Jp (48 bytes) instead of Jc and Jp (128 / 144 bytes): the PCG tail hands over
// vt[c] = G_c p[c] (6 doubles per camera, what K5a gathers instead of p[c]); K5a forms u = Jp (v x X + tau), runs the same per-point
// phase, and folds K(X)^T Jp^T t = (X x g ; g), g = Jp^T t, into the same partial rows; the by-camera gather multiplies each camera's
// row sum by G_c^T once.  Mathematically the same operator; in floating point it differs from the explicit product by rounding
// (1e-16 relative per term) -- and by O(|aa|) <= 1.5e-8 relative for a camera on Ceres' first-order rotation branch with aa != 0
// (there the stored derivative is -[X]x rather than -R [X]x J_r); K4, K7 and the Hessian blocks keep the explicit columns.
__global__ void k_pose_factors(int Nc, const double* __restrict__ camrec, const unsigned char* __restrict__ cam_free, int left_twist,
                               double* __restrict__ G) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= Nc) return;
    double g[36];
#pragma unroll
    for (int k = 0; k < 36; ++k) g[k] = 0.0;
    if (cam_free[c]) {
        const double* cr = camrec + size_t(c) * CAMREC;
        double R[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) R[k] = cr[CR_R + k];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) g[6 * (3 + i) + 3 + j] = R[3 * j + i];            // lower right: R^T
        if (left_twist) {
            const double t0 = cr[CR_T], t1 = cr[CR_T + 1], t2 = cr[CR_T + 2];
            const double s0 = R[0] * t0 + R[3] * t1 + R[6] * t2, s1 = R[1] * t0 + R[4] * t1 + R[7] * t2, s2 = R[2] * t0 + R[5] * t1 + R[8] * t2;   // R^T t
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const double r0 = R[3 * j], r1 = R[3 * j + 1], r2 = R[3 * j + 2];           // column j of R^T
                g[6 * 0 + j] = r0; g[6 * 1 + j] = r1; g[6 * 2 + j] = r2;                    // upper left: R^T
                g[6 * 3 + j] = r1 * s2 - r2 * s1;                                           // lower left: column j of -[s]x R^T = (R^T e_j) x s
                g[6 * 4 + j] = r2 * s0 - r0 * s2;
                g[6 * 5 + j] = r0 * s1 - r1 * s0;
            }
        } else {
            // J_r = I - b [aa]x + cc [aa]x^2,  b = (1 - cos th)/th^2 (K0),  cc = (th - sin th)/th^3 = (1 - a)/th^2
            const double ax = cr[CR_AA], ay = cr[CR_AA + 1], az = cr[CR_AA + 2];
            const double th2 = ax * ax + ay * ay + az * az;
            const double a = cr[CR_A], b = cr[CR_B];
            double cc;
            if (cr[CR_C1] == 0.0) cc = 0.0;          // K0's first-order branch (a = 1, b = 0, c1 = 0): d(RX)/d(aa) = -[X]x, i.e. J_r = I
            else if (th2 < 0.25) cc = 1.0 / 6 + th2 * (-1.0 / 120 + th2 * (1.0 / 5040 + th2 * (-1.0 / 362880 + th2 * (1.0 / 39916800 + th2 * (-1.0 / 6227020800.0 +
                                      th2 * (1.0 / 1307674368000.0 + th2 * (-1.0 / 355687428096000.0)))))));
            else cc = (1.0 - a) / th2;
            const double d = 1.0 - cc * th2;
            g[0] = d + cc * ax * ax;        g[1] = b * az + cc * ax * ay;   g[2] = -b * ay + cc * ax * az;
            g[6] = -b * az + cc * ay * ax;  g[7] = d + cc * ay * ay;        g[8] = b * ax + cc * ay * az;
            g[12] = b * ay + cc * az * ax;  g[13] = -b * ax + cc * az * ay; g[14] = d + cc * az * az;
        }
    }
    double* o = G + 36 * size_t(c);
#pragma unroll
    for (int k = 0; k < 36; ++k) o[k] = g[k];
}

#ifndef BA_K5A_MINB
#define BA_K5A_MINB 6          // min CTAs/SM for ptxas: 6 = 80 registers, no spills
#endif
// scatter_mode 4: one CTA per tile.
template <int MODE, bool FI, int PK>
__global__ void __launch_bounds__(TILE_OBS, BA_K5A_MINB)
k_schur_rows(View V, RowSched S, const double* __restrict__ xin, double* __restrict__ atomic_out, const int* __restrict__ done_flag) {
    __shared__ TileScratch M;
    pdl_trigger();
    const int t = blockIdx.x;
    const int4 ti = __ldg(S.tile_info + t);
    const int o0 = ti.x, n = ti.y & 0xffffff, hbase = ti.z, pfirst = ti.w;
    double yi[4] = {0, 0, 0, 0};
    const int tid = threadIdx.x;
    if (n <= TILE_OBS) {
        // ---- every global load of this tile is issued here, before the first use (one latency, not five) ----
        TileRegs<FI> R;
        R.tm = make_int4(0, 0, -1, 0);
        R.rr = make_double2(0.0, 0.0);
        R.h0 = R.h1 = R.h2 = make_double2(0.0, 0.0);
        if (tid < n) {
            const int o = o0 + tid;
            R.tm = ldg4s(S.tmeta + o);
            const double2* J = V.J2 + o;
            load_jc<PK>(V, J, R.jc);
            load_jp<PK>(V, J, R.jp);
            if (FI) load_ji<PK>(V, J, R.ji);
            if (MODE == SWEEP_RHS) R.rr = V.r2[o];
        }
        if (MODE == SWEEP_MATVEC && (R.tm.w >> 16) > 0) {
            const double2* H = reinterpret_cast<const double2*>(V.Hinv + 6 * size_t(pfirst + tid));
            R.h0 = __ldg(H); R.h1 = __ldg(H + 1); R.h2 = __ldg(H + 2);      // (H0,H1) (H2,H3) (H4,H5)
        }
        // everything above is static during a linear solve; xin and the done flag are the predecessor's (PCG tail) output
        pdl_wait();
        if (done_flag && *done_flag) return;
        tile_rows<MODE, FI>(V, S, M, BarCta{}, tid, n, hbase, pfirst, R, xin, atomic_out, yi);
    } else {
        pdl_wait();
        if (done_flag && *done_flag) return;
        tile_rows_long<MODE, FI, PK>(V, S, M.sred, BarCta{}, tid, o0, n, hbase, pfirst, xin, atomic_out, yi);
    }
    if (FI && V.Ng == 1) tile_intr_store(S, M.sred, BarCta{}, tid, t, yi);
}

// K5a, pose-factored (see k_pose_factors): one CTA per tile, the same schedule, the same per-point phase and camera fold as k_schur_rows;
// streams Jp only.  vt = [G_c p_c per camera, padded to VT_STRIDE doubles | p of the intrinsics entries], pt = the points of the
// linearisation (tile-local rows).
// ncu (profiles/r3c): with 64 bytes per observation this kernel is bound by the L1/shared-memory pipe (93 % busy, DRAM 36 %), i.e. by the
// number of L1 tag requests and shared-memory wavefronts per tile, not by latency or occupancy (12 CTAs/SM and an L2 prefetch of the
// tile two waves ahead were both measured slower than 8 CTAs/SM without).  Hence: vt records padded to 64 bytes and fetched with two
// 32-byte loads (one line per lane instead of three 16-byte requests that straddle lines), the points read per observation straight from
// global memory (lanes of one point share the request), Hinv in registers (or a conflict-free transposed shared-memory copy).
#ifndef BA_K5F_MINB
#define BA_K5F_MINB 8
#endif
#ifndef BA_K5F_XSMEM
#define BA_K5F_XSMEM 0      // 1: the tile's points through shared memory (cp.async by the point threads, one more barrier)
#endif
#ifndef BA_K5F_HSMEM
#define BA_K5F_HSMEM 0      // 1: Hinv through shared memory (transposed, conflict-free) instead of 12 registers
#endif
constexpr int K5F_CTAS_PER_SM = BA_K5F_MINB;
struct TileScratchF {
    alignas(16) double sC[TILE_OBS][6];     // camera-sorted 6-vectors; its first half doubles as sW (per-run point sums) before the second barrier
    double sV[TILE_MAX_PTS][3];
    double sred[8];
#if BA_K5F_HSMEM
    alignas(16) double2 sH[3][TILE_MAX_PTS];
#endif
#if BA_K5F_XSMEM
    double sX[TILE_MAX_PTS][3];
#endif
};
__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))), "l"(gsrc) : "memory");
}

// MODE SWEEP_MATVEC: K5a as described.  MODE SWEEP_RHS: K4 in the same form, t = r - Jp vp (no per-point phase), rows (X x g ; g).
template <int MODE, bool FI, int PK>
__global__ void __launch_bounds__(TILE_OBS, BA_K5F_MINB)
k_schur_rows_f(View V, RowSched S, const double* __restrict__ vt, const double* __restrict__ pt, double* __restrict__ atomic_out,
               const int* __restrict__ done_flag) {
    __shared__ TileScratchF M;
    pdl_trigger();
    const int t = blockIdx.x;
    const int4 ti = __ldg(S.tile_info + t);
    const int o0 = ti.x, n = ti.y & 0xffffff, hbase = ti.z, pfirst = ti.w;
    const int ioff = VT_STRIDE * V.Nc;      // intrinsics entries of vt
    const int yoff = 6 * V.Nc;              // ... and of the output vector
    double yi[4] = {0, 0, 0, 0};
    const int tid = threadIdx.x;
    if (n <= TILE_OBS) {
        // ---- everything static during a linear solve is requested before the dependency wait: schedule record, Jp, the points, Hinv ----
        const bool act = tid < n;
        const int npts = int(unsigned(ti.y) >> 24);
        int4 tm = make_int4(0, 0, -1, 0);
        double2 jp[3], ji[4];
        jp[0] = jp[1] = jp[2] = make_double2(0.0, 0.0);
        double X0 = 0.0, X1 = 0.0, X2 = 0.0;
        double2 h0 = make_double2(0.0, 0.0), h1 = h0, h2 = h0;
        double2 rr = make_double2(0.0, 0.0);
        if (MODE == SWEEP_MATVEC && tid < npts) {   // this thread as a point thread (addressed from tile_info alone: no dependency on the schedule record)
            const double2* H = reinterpret_cast<const double2*>(V.Hinv + 6 * size_t(pfirst + tid));
#if BA_K5F_HSMEM
            cp_async16(&M.sH[0][tid], H); cp_async16(&M.sH[1][tid], H + 1); cp_async16(&M.sH[2][tid], H + 2);
#else
            h0 = __ldg(H); h1 = __ldg(H + 1); h2 = __ldg(H + 2);
#endif
#if BA_K5F_XSMEM
            const double* Xp = pt + 3 * size_t(pfirst + tid);
            cp_async8(&M.sX[tid][0], Xp); cp_async8(&M.sX[tid][1], Xp + 1); cp_async8(&M.sX[tid][2], Xp + 2);
#endif
        }
        if (act) {
            const int o = o0 + tid;
            const uint64_t once = l2_policy_evict_first();
            tm = ldg4_once(S.tmeta + o, once);
            const double2* Jq = V.J2 + o + size_t(6 - (PK < 0 ? V.jpack : PK)) * V.ldo;
            jp[0] = ldg2_once(Jq, once); jp[1] = ldg2_once(Jq + V.ldo, once); jp[2] = ldg2_once(Jq + 2 * size_t(V.ldo), once);
            if (FI) {
#pragma unroll
                for (int k = 0; k < 4; ++k) ji[k] = ldg2_once(Jq + size_t(3 + k) * V.ldo, once);
            }
            if (MODE == SWEEP_RHS) rr = V.r2[o];
#if !BA_K5F_XSMEM
            const double* Xp = pt + 3 * size_t(pfirst + ((tm.y >> 24) & 0xff));
            X0 = __ldg(Xp); X1 = __ldg(Xp + 1); X2 = __ldg(Xp + 2);
#endif
        }
#if BA_K5F_HSMEM || BA_K5F_XSMEM
        cp_async_commit();
        cp_async_wait_all();
        __syncthreads();
#endif
        pdl_wait();
        if (done_flag && *done_flag) return;
        const int npt_obs = tm.w >> 16, pstart = tm.w & 0xffff;   // this thread as a point thread
        const int c = tm.x & 0xffffff;
        const int slot = tm.y & 0xff, rem = (tm.y >> 8) & 0xff, hrank = (tm.y >> 16) & 0xff, lpt = (tm.y >> 24) & 0xff;
        const int g = (FI && act) ? V.cam_group[c] : 0;
        double2 u = make_double2(0.0, 0.0);
        double w0 = 0.0, w1 = 0.0, w2 = 0.0;
        double tx = 0.0, ty = 0.0;
      if (MODE == SWEEP_MATVEC) {
        if (act) {
            double x[8];                               // v = x[0..2], tau = x[3..5]
#if BA_K5F_VT256
            ldg256(vt + VT_STRIDE * size_t(c), x); ldg256(vt + VT_STRIDE * size_t(c) + 4, x + 4);
#else
            {
                const double2* xc = reinterpret_cast<const double2*>(vt + VT_STRIDE * size_t(c));
                const double2 q0 = __ldg(xc), q1 = __ldg(xc + 1), q2 = __ldg(xc + 2);
                x[0] = q0.x; x[1] = q0.y; x[2] = q1.x; x[3] = q1.y; x[4] = q2.x; x[5] = q2.y;
            }
#endif
#if BA_K5F_XSMEM
            X0 = M.sX[lpt][0]; X1 = M.sX[lpt][1]; X2 = M.sX[lpt][2];
#endif
            const double d0 = x[1] * X2 - x[2] * X1 + x[3];
            const double d1 = x[2] * X0 - x[0] * X2 + x[4];
            const double d2 = x[0] * X1 - x[1] * X0 + x[5];
            u.x = jp[0].x * d0 + jp[1].x * d1 + jp[2].x * d2;
            u.y = jp[0].y * d0 + jp[1].y * d1 + jp[2].y * d2;
            if (FI) {
#pragma unroll
                for (int k = 0; k < 4; ++k) { const double xv = __ldg(vt + ioff + 4 * g + k); u.x += ji[k].x * xv; u.y += ji[k].y * xv; }
            }
            w0 = jp[0].x * u.x + jp[0].y * u.y; w1 = jp[1].x * u.x + jp[1].y * u.y; w2 = jp[2].x * u.x + jp[2].y * u.y;
        }
        double (*sW)[3] = reinterpret_cast<double (*)[3]>(&M.sC[0][0]);
        {   // per-point sums: segmented shuffle reduction inside the warp, one shared-memory value per run (as in tile_rows)
            const unsigned same = __match_any_sync(0xffffffffu, act ? lpt : 0xffff);
            const int lane = tid & 31;
            const int segrem = __popc(same >> lane) - 1;
            const int maxseg = __reduce_max_sync(0xffffffffu, segrem);
            for (int d = 1; d <= maxseg; d <<= 1) {
                const double a0 = __shfl_down_sync(0xffffffffu, w0, d), a1 = __shfl_down_sync(0xffffffffu, w1, d), a2 = __shfl_down_sync(0xffffffffu, w2, d);
                if (d <= segrem) { w0 += a0; w1 += a1; w2 += a2; }
            }
            if (act && (same & ((1u << lane) - 1u)) == 0) { sW[tid][0] = w0; sW[tid][1] = w1; sW[tid][2] = w2; }
        }
        __syncthreads();
        if (npt_obs > 0) {
            double W0 = sW[pstart][0], W1 = sW[pstart][1], W2 = sW[pstart][2];
            for (int i = (pstart / 32 + 1) * 32, e = pstart + npt_obs; i < e; i += 32) { W0 += sW[i][0]; W1 += sW[i][1]; W2 += sW[i][2]; }
#if BA_K5F_HSMEM
            h0 = M.sH[0][tid]; h1 = M.sH[1][tid]; h2 = M.sH[2][tid];
#endif
            // (H0,H1) (H2,H3) (H4,H5); Hinv rows (H0,H1,H2) (H1,H3,H4) (H2,H4,H5)
            M.sV[tid][0] = h0.x * W0 + h0.y * W1 + h1.x * W2;
            M.sV[tid][1] = h0.y * W0 + h1.y * W1 + h2.x * W2;
            M.sV[tid][2] = h1.x * W0 + h2.x * W1 + h2.y * W2;
        }
        __syncthreads();
        if (act) {
            const double* v = M.sV[lpt];
            tx = u.x - (jp[0].x * v[0] + jp[1].x * v[1] + jp[2].x * v[2]);
            ty = u.y - (jp[0].y * v[0] + jp[1].y * v[1] + jp[2].y * v[2]);
        }
      } else if (act) {   // SWEEP_RHS: t = r - E vp
            const double* v = V.vp + 3 * size_t(pfirst + lpt);
            tx = rr.x - (jp[0].x * v[0] + jp[1].x * v[1] + jp[2].x * v[2]);
            ty = rr.y - (jp[0].y * v[0] + jp[1].y * v[1] + jp[2].y * v[2]);
      }
        if (act) {
            const double g0 = jp[0].x * tx + jp[0].y * ty, g1 = jp[1].x * tx + jp[1].y * ty, g2 = jp[2].x * tx + jp[2].y * ty;
#if BA_K5F_XSMEM
            X0 = M.sX[lpt][0]; X1 = M.sX[lpt][1]; X2 = M.sX[lpt][2];
#endif
            double2* sc = reinterpret_cast<double2*>(M.sC[slot]);
            sc[0] = make_double2(X1 * g2 - X2 * g1, X2 * g0 - X0 * g2);
            sc[1] = make_double2(X0 * g1 - X1 * g0, g0);
            sc[2] = make_double2(g1, g2);
            if (FI) {
#pragma unroll
                for (int k = 0; k < 4; ++k) yi[k] = ji[k].x * tx + ji[k].y * ty;
                if (V.Ng > 1) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) red_add(atomic_out + yoff + 4 * g + k, yi[k]);
                }
            }
        }
        __syncthreads();
        {   // camera fold over 32 sorted positions (as in tile_rows)
            double2 a = make_double2(0.0, 0.0), bq = a, cq = a;
            if (act) {
                const double2* row = reinterpret_cast<const double2*>(M.sC[tid]);
                a = row[0]; bq = row[1]; cq = row[2];
            }
            // shuffles share the shared-memory data pipe, the unit this kernel saturates: only as many steps as the longest run of this warp needs
            const int maxrem = __reduce_max_sync(0xffffffffu, act ? rem : 0);
            for (int d = 1; d <= maxrem; d <<= 1) {
                const double ax = __shfl_down_sync(0xffffffffu, a.x, d), ay = __shfl_down_sync(0xffffffffu, a.y, d);
                const double bx = __shfl_down_sync(0xffffffffu, bq.x, d), by = __shfl_down_sync(0xffffffffu, bq.y, d);
                const double cx = __shfl_down_sync(0xffffffffu, cq.x, d), cy = __shfl_down_sync(0xffffffffu, cq.y, d);
                if (d <= rem) { a.x += ax; a.y += ay; bq.x += bx; bq.y += by; cq.x += cx; cq.y += cy; }
            }
            if (tm.z >= 0) {
                double2* dst = reinterpret_cast<double2*>(S.rows + size_t(ROW_DOUBLES) * (hbase + hrank));
                const uint64_t keep = l2_policy_evict_last();
                row_store_keep(dst, a, keep); row_store_keep(dst + 1, bq, keep); row_store_keep(dst + 2, cq, keep);
            }
        }
    } else {
        // one point with more than TILE_OBS observations: two passes from global memory, every observation its own row
        pdl_wait();
        if (done_flag && *done_flag) return;
        const int o1 = o0 + n;
        const double X0 = pt[3 * size_t(pfirst)], X1 = pt[3 * size_t(pfirst) + 1], X2 = pt[3 * size_t(pfirst) + 2];
        auto forward = [&](int o, int c, const double2 (&jp)[3]) {
            const double* x = vt + VT_STRIDE * size_t(c);
            const double d0 = x[1] * X2 - x[2] * X1 + x[3], d1 = x[2] * X0 - x[0] * X2 + x[4], d2 = x[0] * X1 - x[1] * X0 + x[5];
            double2 u = make_double2(jp[0].x * d0 + jp[1].x * d1 + jp[2].x * d2, jp[0].y * d0 + jp[1].y * d1 + jp[2].y * d2);
            if (FI) {
                const int g = V.cam_group[c];
                double2 ji[4];
                load_ji<PK>(V, V.J2 + o, ji);
#pragma unroll
                for (int k = 0; k < 4; ++k) { const double xv = vt[ioff + 4 * g + k]; u.x += ji[k].x * xv; u.y += ji[k].y * xv; }
            }
            return u;
        };
        double v0, v1, v2;
        if (MODE == SWEEP_MATVEC) {
            double W[3] = {0, 0, 0};
            for (int o = o0 + tid; o < o1; o += TILE_OBS) {
                double2 jp[3];
                load_jp<PK>(V, V.J2 + o, jp);
                const double2 u = forward(o, V.obs_cam[o], jp);
#pragma unroll
                for (int a = 0; a < 3; ++a) W[a] += jp[a].x * u.x + jp[a].y * u.y;
            }
            double w[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) w[a] = tile_sum(W[a], M.sred, BarCta{}, tid);
            const double* H = V.Hinv + 6 * size_t(pfirst);
            v0 = H[0] * w[0] + H[1] * w[1] + H[2] * w[2]; v1 = H[1] * w[0] + H[3] * w[1] + H[4] * w[2]; v2 = H[2] * w[0] + H[4] * w[1] + H[5] * w[2];
        } else { v0 = V.vp[3 * size_t(pfirst)]; v1 = V.vp[3 * size_t(pfirst) + 1]; v2 = V.vp[3 * size_t(pfirst) + 2]; }
        for (int o = o0 + tid; o < o1; o += TILE_OBS) {
            const int c = V.obs_cam[o];
            double2 jp[3];
            load_jp<PK>(V, V.J2 + o, jp);
            const double2 u = MODE == SWEEP_MATVEC ? forward(o, c, jp) : V.r2[o];
            const double tx = u.x - (jp[0].x * v0 + jp[1].x * v1 + jp[2].x * v2), ty = u.y - (jp[0].y * v0 + jp[1].y * v1 + jp[2].y * v2);
            const double g0 = jp[0].x * tx + jp[0].y * ty, g1 = jp[1].x * tx + jp[1].y * ty, g2 = jp[2].x * tx + jp[2].y * ty;
            double2* dst = reinterpret_cast<double2*>(S.rows + size_t(ROW_DOUBLES) * (hbase + (o - o0)));
            row_store(dst, make_double2(X1 * g2 - X2 * g1, X2 * g0 - X0 * g2));
            row_store(dst + 1, make_double2(X0 * g1 - X1 * g0, g0));
            row_store(dst + 2, make_double2(g1, g2));
            if (FI) {
                const int g = V.cam_group[c];
                double2 ji[4];
                load_ji<PK>(V, V.J2 + o, ji);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const double contrib = ji[k].x * tx + ji[k].y * ty;
                    if (V.Ng > 1) red_add(atomic_out + yoff + 4 * g + k, contrib); else yi[k] += contrib;
                }
            }
        }
    }
    if (FI && V.Ng == 1) tile_intr_store(S, M.sred, BarCta{}, tid, t, yi);
}

// pt_by_cam[i] = local point of the i-th observation in camera order (static; built on the device once per ba_set_problem)
__global__ void k_points_by_camera(int No, const int* __restrict__ perm_by_cam, const int* __restrict__ obs_pt, int* __restrict__ pt_by_cam) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < No) pt_by_cam[i] = obs_pt[perm_by_cam[i]];
}

// K2c, pose-factored: Hcc[c] = G^T N G and gc[c] = G^T n with  N = sum K(X)^T (Jp^T Jp) K(X),  n = sum K(X)^T Jp^T r  over the camera's
// observations (camera order, fixed summation order => deterministic), K(X) = [-[X]x | I].  Reads the SoA point columns and residuals K1
// writes for K4/K5/K7 anyway (four 16-byte pieces per observation) and the point, so K1 need not write the 128-byte by-observation
// records (640 MB per evaluation on the Venice shape).  With free intrinsics also the camera's share of its group's 4x4 block and
// gradient (ipart, as k2c_camera_blocks<true>).  N = [ -S M S   S M ; (S M)^T   M ],  S = [X]x,  M = Jp^T Jp.
#ifndef BA_K2CF_MINB
#define BA_K2CF_MINB 4
#endif
template <bool FI, int PK>
__global__ void __launch_bounds__(128, FI ? 3 : BA_K2CF_MINB)
k2c_camera_blocks_f(View V, const double* __restrict__ G, const double* __restrict__ pt, const int* __restrict__ pt_by_cam,
                    double* __restrict__ Hcc, double* __restrict__ gc, double* __restrict__ ipart) {
    constexpr int NV = FI ? (21 + 6 + 10 + 4) : (21 + 6);   // N_rr (6), S M (9), M (6), n (6), [Hii upper, gi]
    __shared__ double sm[4][NV];
    __shared__ double sN[36], sT[36], sG[36], sn[6];
    const int c = blockIdx.x;
    const int a0 = V.cam_off[c], a1 = V.cam_off[c + 1];
    if (threadIdx.x < 36) sG[threadIdx.x] = G[36 * size_t(c) + threadIdx.x];
    const size_t ld = V.ldo;
    const int jp_col = 6 - (PK < 0 ? V.jpack : PK);
    double acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k] = 0.0;
    for (int i = a0 + threadIdx.x; i < a1; i += 128) {
        const int o = __ldg(V.perm_by_cam + i), p = __ldg(pt_by_cam + i);
        const double2* Jq = V.J2 + o + size_t(jp_col) * ld;
        const double2 j0 = __ldg(Jq), j1 = __ldg(Jq + ld), j2 = __ldg(Jq + 2 * ld), r = __ldg(V.r2 + o);
        const double X0 = __ldg(pt + 3 * size_t(p)), X1 = __ldg(pt + 3 * size_t(p) + 1), X2 = __ldg(pt + 3 * size_t(p) + 2);
        const double M00 = j0.x * j0.x + j0.y * j0.y, M01 = j0.x * j1.x + j0.y * j1.y, M02 = j0.x * j2.x + j0.y * j2.y;
        const double M11 = j1.x * j1.x + j1.y * j1.y, M12 = j1.x * j2.x + j1.y * j2.y, M22 = j2.x * j2.x + j2.y * j2.y;
        // P = S M
        const double P00 = X1 * M02 - X2 * M01, P01 = X1 * M12 - X2 * M11, P02 = X1 * M22 - X2 * M12;
        const double P10 = X2 * M00 - X0 * M02, P11 = X2 * M01 - X0 * M12, P12 = X2 * M02 - X0 * M22;
        const double P20 = X0 * M01 - X1 * M00, P21 = X0 * M11 - X1 * M01, P22 = X0 * M12 - X1 * M02;
        // N_rr = -P S (symmetric): (P S)_i0 = P_i1 X2 - P_i2 X1, (P S)_i1 = P_i2 X0 - P_i0 X2, (P S)_i2 = P_i0 X1 - P_i1 X0
        acc[0] += P02 * X1 - P01 * X2;  acc[1] += P00 * X2 - P02 * X0;  acc[2] += P01 * X0 - P00 * X1;
        acc[3] += P10 * X2 - P12 * X0;  acc[4] += P11 * X0 - P10 * X1;  acc[5] += P21 * X0 - P20 * X1;
        acc[6] += P00; acc[7] += P01; acc[8] += P02; acc[9] += P10; acc[10] += P11; acc[11] += P12; acc[12] += P20; acc[13] += P21; acc[14] += P22;
        acc[15] += M00; acc[16] += M01; acc[17] += M02; acc[18] += M11; acc[19] += M12; acc[20] += M22;
        const double g0 = j0.x * r.x + j0.y * r.y, g1 = j1.x * r.x + j1.y * r.y, g2 = j2.x * r.x + j2.y * r.y;
        acc[21] += X1 * g2 - X2 * g1; acc[22] += X2 * g0 - X0 * g2; acc[23] += X0 * g1 - X1 * g0;
        acc[24] += g0; acc[25] += g1; acc[26] += g2;
        if (FI) {
            double2 ji[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) ji[k] = __ldg(Jq + size_t(3 + k) * ld);
            int n = 27;
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = a; b < 4; ++b) acc[n++] += ji[a].x * ji[b].x + ji[a].y * ji[b].y;
#pragma unroll
            for (int a = 0; a < 4; ++a) acc[n++] += ji[a].x * r.x + ji[a].y * r.y;
        }
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NV; ++k) { const double s = warp_sum(acc[k]); if (lane == 0) sm[wid][k] = s; }
    __syncthreads();
    if (threadIdx.x < NV) {
        const int k = threadIdx.x;
        const double s = (sm[0][k] + sm[1][k]) + (sm[2][k] + sm[3][k]);
        if (k < 6) {             // N_rr upper triangle: (0,0) (0,1) (0,2) (1,1) (1,2) (2,2)
            const int a = k < 3 ? 0 : (k < 5 ? 1 : 2), b = k < 3 ? k : (k < 5 ? k - 2 : 2);
            sN[6 * a + b] = s; sN[6 * b + a] = s;
        } else if (k < 15) {     // N_rt = S M (3x3) and its transpose
            const int a = (k - 6) / 3, b = (k - 6) % 3;
            sN[6 * a + 3 + b] = s; sN[6 * (3 + b) + a] = s;
        } else if (k < 21) {     // N_tt = M upper triangle
            const int q = k - 15, a = q < 3 ? 0 : (q < 5 ? 1 : 2), b = q < 3 ? q : (q < 5 ? q - 2 : 2);
            sN[6 * (3 + a) + 3 + b] = s; sN[6 * (3 + b) + 3 + a] = s;
        } else if (k < 27) sn[k - 21] = s;
        else if (FI) ipart[14 * size_t(c) + (k - 27)] = s;
    }
    __syncthreads();
    if (threadIdx.x < 36) {      // T = N G
        const int a = threadIdx.x / 6, b = threadIdx.x % 6;
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < 6; ++k) t += sN[6 * a + k] * sG[6 * k + b];
        sT[threadIdx.x] = t;
    }
    __syncthreads();
    if (threadIdx.x < 36) {      // Hcc = G^T T; the upper triangle is computed and mirrored so that the block is exactly symmetric
        const int a = threadIdx.x / 6, b = threadIdx.x % 6;
        const int lo = min(a, b), hi = max(a, b);
        double h = 0.0;
#pragma unroll
        for (int k = 0; k < 6; ++k) h += sG[6 * k + lo] * sT[6 * k + hi];
        Hcc[36 * size_t(c) + threadIdx.x] = h;
    } else if (threadIdx.x < 42) {
        const int a = threadIdx.x - 36;
        double v = 0.0;
#pragma unroll
        for (int k = 0; k < 6; ++k) v += sG[6 * k + a] * sn[k];
        gc[6 * size_t(c) + a] = v;
    }
}

// vt = [G_c x_c | intrinsics entries of x] for an arbitrary reduced vector x (K7's input: the PCG solution).
__global__ void k_pose_apply(int Nc, int nf, const double* __restrict__ G, const double* __restrict__ x, double* __restrict__ vt) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nf) return;
    if (j < 6 * Nc) {
        const int c = j / 6, a = j - 6 * c;
        const double* Gr = G + 36 * size_t(c) + 6 * a;
        const double* xc = x + 6 * c;
        double v = 0.0;
#pragma unroll
        for (int k = 0; k < 6; ++k) v += Gr[k] * xc[k];
        vt[VT_STRIDE * size_t(c) + a] = v;
    } else vt[j + (VT_STRIDE - 6) * Nc] = x[j];
}

// K7, pose-factored: yp = vp - Hinv E^T F y and the partial of sum m.(r + m/2), m = -(F y + E yp), with F y = Jp (v x X + tau) [+ Ji y_i],
// (v, tau) = G_c y_c from k_pose_apply.  Same structure as k_schur_sweep<SWEEP_BACKSUB>; streams Jp, r and the index arrays only.
template <bool FI, int PK>
__global__ void __launch_bounds__(TILE_OBS)
k_backsub_f(View V, const double* __restrict__ vt, const double* __restrict__ pt, double* __restrict__ partial) {
    __shared__ double sW[TILE_OBS][3];
    __shared__ double sV[TILE_OBS][3];
    __shared__ double sred[32];
    const int t = blockIdx.x;
    const int o0 = V.tile_off[t], o1 = V.tile_off[t + 1];
    const int ioff = VT_STRIDE * V.Nc;
    double acc = 0.0;
    auto forward = [&](int o, int c, const double* X, const double2 (&jp)[3]) {
        const double2* xc = reinterpret_cast<const double2*>(vt + VT_STRIDE * size_t(c));
        const double2 q0 = __ldg(xc), q1 = __ldg(xc + 1), q2 = __ldg(xc + 2);
        const double d0 = q0.y * X[2] - q1.x * X[1] + q1.y, d1 = q1.x * X[0] - q0.x * X[2] + q2.x, d2 = q0.x * X[1] - q0.y * X[0] + q2.y;
        double2 u = make_double2(jp[0].x * d0 + jp[1].x * d1 + jp[2].x * d2, jp[0].y * d0 + jp[1].y * d1 + jp[2].y * d2);
        if (FI) {
            const int g = V.cam_group[c];
            double2 ji[4];
            load_ji<PK>(V, V.J2 + o, ji);
#pragma unroll
            for (int k = 0; k < 4; ++k) { const double xv = __ldg(vt + ioff + 4 * g + k); u.x += ji[k].x * xv; u.y += ji[k].y * xv; }
        }
        return u;
    };
    if (o1 - o0 <= TILE_OBS) {
        const int o = o0 + threadIdx.x;
        const bool act = o < o1;
        double2 jp[3];
        double2 u = make_double2(0.0, 0.0);
        int p = 0;
        if (act) {
            const int c = V.obs_cam[o];
            p = V.obs_pt[o];
            load_jp<PK>(V, V.J2 + o, jp);
            const double X[3] = {__ldg(pt + 3 * size_t(p)), __ldg(pt + 3 * size_t(p) + 1), __ldg(pt + 3 * size_t(p) + 2)};
            u = forward(o, c, X, jp);
        }
        {   // per-point sums: segmented shuffle reduction inside the warp (a point's observations are adjacent lanes), one shared-memory
            // value per run; the point thread then adds one value per warp its point touches (as K5a does)
            double w0 = 0.0, w1 = 0.0, w2 = 0.0;
            if (act) { w0 = jp[0].x * u.x + jp[0].y * u.y; w1 = jp[1].x * u.x + jp[1].y * u.y; w2 = jp[2].x * u.x + jp[2].y * u.y; }
            const unsigned same = __match_any_sync(0xffffffffu, act ? p : -1);
            const int lane = threadIdx.x & 31;
            const int segrem = __popc(same >> lane) - 1;
            const int maxseg = __reduce_max_sync(0xffffffffu, segrem);
            for (int d = 1; d <= maxseg; d <<= 1) {
                const double a0 = __shfl_down_sync(0xffffffffu, w0, d), a1 = __shfl_down_sync(0xffffffffu, w1, d), a2 = __shfl_down_sync(0xffffffffu, w2, d);
                if (d <= segrem) { w0 += a0; w1 += a1; w2 += a2; }
            }
            if (act && (same & ((1u << lane) - 1u)) == 0) { sW[threadIdx.x][0] = w0; sW[threadIdx.x][1] = w1; sW[threadIdx.x][2] = w2; }
        }
        __syncthreads();
        const int pfirst = V.obs_pt[o0], plast = V.obs_pt[o1 - 1];
        const int pj = pfirst + threadIdx.x;
        if (pj <= plast) {
            const int i0 = V.pt_off[pj] - o0, i1 = V.pt_off[pj + 1] - o0;
            double W[3] = {0, 0, 0};
            if (i1 > i0) {
                W[0] = sW[i0][0]; W[1] = sW[i0][1]; W[2] = sW[i0][2];
                for (int i = (i0 / 32 + 1) * 32; i < i1; i += 32) { W[0] += sW[i][0]; W[1] += sW[i][1]; W[2] += sW[i][2]; }
            }
            const double* H = V.Hinv + 6 * size_t(pj);
            const double v0 = V.vp[3 * size_t(pj)] - (H[0] * W[0] + H[1] * W[1] + H[2] * W[2]);
            const double v1 = V.vp[3 * size_t(pj) + 1] - (H[1] * W[0] + H[3] * W[1] + H[4] * W[2]);
            const double v2 = V.vp[3 * size_t(pj) + 2] - (H[2] * W[0] + H[4] * W[1] + H[5] * W[2]);
            V.yp[3 * size_t(pj)] = v0; V.yp[3 * size_t(pj) + 1] = v1; V.yp[3 * size_t(pj) + 2] = v2;
            sV[threadIdx.x][0] = v0; sV[threadIdx.x][1] = v1; sV[threadIdx.x][2] = v2;
        }
        __syncthreads();
        if (act) {
            const double* v = sV[p - pfirst];
            const double e0 = jp[0].x * v[0] + jp[1].x * v[1] + jp[2].x * v[2];
            const double e1 = jp[0].y * v[0] + jp[1].y * v[1] + jp[2].y * v[2];
            const double m0 = -(u.x + e0), m1 = -(u.y + e1);
            const double2 r = V.r2[o];
            acc += m0 * (r.x + 0.5 * m0) + m1 * (r.y + 0.5 * m1);
        }
    } else {   // one point with more than TILE_OBS observations
        const int p = V.obs_pt[o0];
        const double X[3] = {pt[3 * size_t(p)], pt[3 * size_t(p) + 1], pt[3 * size_t(p) + 2]};
        double W[3] = {0, 0, 0};
        for (int o = o0 + threadIdx.x; o < o1; o += TILE_OBS) {
            double2 jp[3];
            load_jp<PK>(V, V.J2 + o, jp);
            const double2 u = forward(o, V.obs_cam[o], X, jp);
#pragma unroll
            for (int a = 0; a < 3; ++a) W[a] += jp[a].x * u.x + jp[a].y * u.y;
        }
#pragma unroll
        for (int a = 0; a < 3; ++a) { const double s = block_sum(W[a], sred); if (threadIdx.x == 0) sV[0][a] = s; }
        __syncthreads();
        if (threadIdx.x == 0) {
            const double* H = V.Hinv + 6 * size_t(p);
            const double w0 = sV[0][0], w1 = sV[0][1], w2 = sV[0][2];
            const double v0 = V.vp[3 * size_t(p)] - (H[0] * w0 + H[1] * w1 + H[2] * w2);
            const double v1 = V.vp[3 * size_t(p) + 1] - (H[1] * w0 + H[3] * w1 + H[4] * w2);
            const double v2 = V.vp[3 * size_t(p) + 2] - (H[2] * w0 + H[4] * w1 + H[5] * w2);
            V.yp[3 * size_t(p)] = v0; V.yp[3 * size_t(p) + 1] = v1; V.yp[3 * size_t(p) + 2] = v2;
            sV[1][0] = v0; sV[1][1] = v1; sV[1][2] = v2;
        }
        __syncthreads();
        const double v0 = sV[1][0], v1 = sV[1][1], v2 = sV[1][2];
        for (int o = o0 + threadIdx.x; o < o1; o += TILE_OBS) {
            double2 jp[3];
            load_jp<PK>(V, V.J2 + o, jp);
            const double2 u = forward(o, V.obs_cam[o], X, jp);
            const double e0 = jp[0].x * v0 + jp[1].x * v1 + jp[2].x * v2, e1 = jp[0].y * v0 + jp[1].y * v1 + jp[2].y * v2;
            const double m0 = -(u.x + e0), m1 = -(u.y + e1);
            const double2 r = V.r2[o];
            acc += m0 * (r.x + 0.5 * m0) + m1 * (r.y + 0.5 * m1);
        }
    }
    const double s = block_sum(acc, sred);
    if (threadIdx.x == 0) partial[t] = s;
}

// ---------------------------------------------------------------------------------------------------------
// K5b — y[c] = sum of camera c's partial rows: one 4-warp CTA per camera, fixed combine order (bit-reproducible).
// Four adjacent lanes share one row (lane k loads its k-th 16 bytes), so a warp instruction touches 8 rows = 8 lines; the first
// version gave every lane a whole row (3 L1 wavefronts per row, the kernel throttled by the load/store unit).
// The row list is static during a linear solve: the CTA copies (up to GATHER_IDX_CAP entries of) it into shared memory BEFORE
// griddepcontrol.wait, i.e. while K5a is still running; after the wait there is no index -> row dependency left, only row loads,
// GATHER_UNR of them in flight per lane.
#ifndef BA_GATHER_UNR
#define BA_GATHER_UNR 4
#endif
#ifndef BA_GATHER_MINB
#define BA_GATHER_MINB 16      // min CTAs/SM (32 registers): all 1779 CTAs of the Venice shape are ONE wave (148 x 16 = 2368).  With 12 per SM
                               // (1776 slots) three CTAs had to wait for a first-wave CTA to finish, and the PCG tail -- a programmatic dependent
                               // launch that needs every gather CTA started -- waited with them: ~45 us per PCG iteration behind the streaming
                               // kernel, whose CTAs leave no room for early gather CTAs.
#endif
constexpr int GATHER_IDX_CAP = 1536;   // row ids staged per CTA (6 KB); longer lists read the remainder from global memory

// Whole CTA.  On return (after a CTA barrier) sw[0..5] holds the 6-vector.  Returns false (whole CTA, before any barrier) when *done
// says the PCG loop has already stopped.
__device__ __forceinline__ bool camera_row_sum(const RowSched& S, int c, double* sw, int* sidx, const int* done) {
    constexpr int NG = GATHER_WARPS * 8;          // 4-lane groups per CTA
    constexpr int UNR = BA_GATHER_UNR;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = __ldg(S.cam_row_off + c), e = __ldg(S.cam_row_off + c + 1);
    const int cnt = e - b, staged = min(cnt, GATHER_IDX_CAP);
    for (int i = threadIdx.x; i < staged; i += GATHER_WARPS * 32) sidx[i] = __ldg(S.cam_rows + b + i);
    __syncthreads();
    pdl_wait();
    if (done && *done) return false;
    const int sub = threadIdx.x & 3;
    const double2* R2 = reinterpret_cast<const double2*>(S.rows);
    double2 acc = make_double2(0.0, 0.0);
    if (sub < 3) {
        const uint64_t keep = l2_policy_evict_last();
        int i = threadIdx.x >> 2;
        for (; i + (UNR - 1) * NG < staged; i += UNR * NG) {     // UNR rows per group and trip, all in flight
            double2 v[UNR];
#pragma unroll
            for (int k = 0; k < UNR; ++k) v[k] = row_load_keep(R2 + size_t(ROW_DOUBLES / 2) * sidx[i + k * NG] + sub, keep);
#pragma unroll
            for (int k = 1; k < UNR; ++k) { v[0].x += v[k].x; v[0].y += v[k].y; }
            acc.x += v[0].x; acc.y += v[0].y;
        }
        for (; i < staged; i += NG) { const double2 v = row_load_keep(R2 + size_t(ROW_DOUBLES / 2) * sidx[i] + sub, keep); acc.x += v.x; acc.y += v.y; }
        for (; i < cnt; i += NG) { const double2 v = row_load_keep(R2 + size_t(ROW_DOUBLES / 2) * __ldg(S.cam_rows + b + i) + sub, keep); acc.x += v.x; acc.y += v.y; }
    }
    // the 8 groups of a warp (lanes with equal sub), fixed tree; then the warps in order
#pragma unroll
    for (int off = 4; off < 32; off <<= 1) { acc.x += __shfl_xor_sync(0xffffffffu, acc.x, off); acc.y += __shfl_xor_sync(0xffffffffu, acc.y, off); }
    __shared__ double swp[GATHER_WARPS][6];
    if (lane < 3) { swp[wid][2 * lane] = acc.x; swp[wid][2 * lane + 1] = acc.y; }
    __syncthreads();
    if (threadIdx.x < 6) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < GATHER_WARPS; ++w) v += swp[w][threadIdx.x];
        sw[threadIdx.x] = v;
    }
    __syncthreads();
    return true;
}

// The intrinsics entries of the matvec (CTA `Nc` of the gather kernels): single group -> ordered sum of the per-tile partials; several
// groups -> RED-accumulated in yin already.  f(j, v) stores entry j.
template <class F>
__device__ __forceinline__ void gather_intrinsics(int Nc, int Ng, int nTiles, const RowSched& S, const double* yin, F&& f) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int k = wid; k < 4 * Ng; k += GATHER_WARPS) {
        double v;
        if (Ng == 1) {
            double s = 0.0;
            for (int t = lane; t < nTiles; t += 32) s += __ldcg(S.tile_intr + 4 * size_t(t) + k);
            v = warp_sum(s);
        } else v = yin[6 * Nc + k];
        if (lane == 0) f(6 * Nc + k, v);
    }
}

// Entry j of G^T w (G row-major 6x6 in shared memory), fixed order.
__device__ __forceinline__ double pose_factor_apply_t(const double* sG, const double* w, int j) {
    return ((sG[j] * w[0] + sG[6 + j] * w[1]) + (sG[12 + j] * w[2] + sG[18 + j] * w[3])) + (sG[24 + j] * w[4] + sG[30 + j] * w[5]);
}

// qdot = 0: y[c] = sum of rows (K4: the reduced right-hand side; multi-GPU NCCL path).  G != nullptr: the rows are pose-factored
// (k_schur_rows_f) and the sum is multiplied by G_c^T.
// qdot = 1 (one GPU, PCG loop): q = y + Df2 p stored in y, and the CTA's share of p.q in dotp[blockIdx.x] — the PCG tail adds the
//          Nc + 1 shares in index order (no ticket, no atomics).
__global__ void __launch_bounds__(GATHER_WARPS * 32, BA_GATHER_MINB)
k_camera_gather(int Nc, int Ng, int fi, int nTiles, RowSched S, const double* __restrict__ Df2, const double* __restrict__ p,
                double* __restrict__ y, double* __restrict__ dotp, const PcgState2* __restrict__ st, int qdot, const double* __restrict__ G) {
    __shared__ double sw[6];
    __shared__ int sidx[GATHER_IDX_CAP];
    __shared__ double sdot[GATHER_WARPS];
    __shared__ double sG[36];
    pdl_trigger();
    const int c = blockIdx.x;
    if (c < Nc) {
        if (G && threadIdx.x < 36) sG[threadIdx.x] = __ldg(G + 36 * size_t(c) + threadIdx.x);   // static during the solve; visible after camera_row_sum's barriers
        if (!camera_row_sum(S, c, sw, sidx, st ? &st->done : nullptr)) return;
        if (threadIdx.x < 6) {
            double v = sw[threadIdx.x];
            if (G) v = pose_factor_apply_t(sG, sw, threadIdx.x);   // pose-factored rows: y_c = G_c^T (sum of rows)
            const int j = 6 * c + threadIdx.x;
            double dot = 0.0;
            // p is the predecessor grid's output and this kernel's CTAs may share an SM with a successor's (the next K5a is launched while
            // this grid still runs): read it past L1 (ld.global.cg) so that no L1 line of p exists that the PCG tail's update could leave stale
            if (qdot) { const double pj = __ldcg(p + j); v += Df2[j] * pj; dot = pj * v; }
            y[j] = v;
            if (qdot) {   // lanes 0..5 of warp 0, fixed order
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < 6; ++k) s += __shfl_sync(0x3fu, dot, k, 8);
                if (threadIdx.x == 0) dotp[c] = s;
            }
        }
    } else {
        pdl_wait();
        if (st && st->done) return;
        double dot = 0.0;
        if (fi) gather_intrinsics(Nc, Ng, nTiles, S, y, [&](int j, double v) {
            if (qdot) { const double pj = __ldcg(p + j); v += Df2[j] * pj; dot += pj * v; }
            y[j] = v;
        });
        if (qdot) {   // lane 0 of each warp holds its entries' share
            if ((threadIdx.x & 31) == 0) sdot[threadIdx.x >> 5] = dot;
            __syncthreads();
            if (threadIdx.x == 0) { double s = 0.0; for (int w = 0; w < GATHER_WARPS; ++w) s += sdot[w]; dotp[Nc] = s; }
        }
    }
}

}  // namespace bak
