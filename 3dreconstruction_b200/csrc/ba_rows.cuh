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

#ifndef BA_ROW_DOUBLES
#define BA_ROW_DOUBLES 6
#endif
constexpr int ROW_DOUBLES = BA_ROW_DOUBLES;    // 6 used; 8: 64-byte rows = two full 32-byte sectors; 6: packed 48-byte rows (less to write, the gather still fetches two sectors per row)
// Cache operators of the partial rows (A/B knobs): store 0 = st.cg, 1 = st.cs (streaming), 2 = default; load 0 = ld.cg, 1 = ld.cs, 2 = ld.lu (last use)
#ifndef BA_ROW_ST
#define BA_ROW_ST 0
#endif
#ifndef BA_ROW_LD
#define BA_ROW_LD 0
#endif
__device__ __forceinline__ void row_store(double2* p, double2 v) {
#if BA_ROW_ST == 1
    __stcs(p, v);
#elif BA_ROW_ST == 2
    *p = v;
#else
    __stcg(p, v);
#endif
}
__device__ __forceinline__ double2 row_load(const double2* p) {
#if BA_ROW_LD == 1
    return __ldcs(p);
#elif BA_ROW_LD == 2
    return __ldlu(p);
#else
    return __ldcg(p);
#endif
}
constexpr int L1_CHUNK = 8;       // level-1 chunk of the in-tile segmented sum
constexpr int GATHER_WARPS = 4;   // warps per CTA of the by-camera gather (K5b)

struct RowSched {
    // One 16-byte record per tile position i = o0 + tid, read with a single LDG.128 at kernel entry:
    //   .x  camera of observation i | sorted position of the head of its camera segment << 24
    //   .y  slot | lvl1 << 8 | hrank << 16 | lpt << 24   (observation i: rank in the tile's camera-sorted order; sorted position
    //       tid: rows to fold in level 1, index of this segment head among the tile's heads; observation i: tile-local point)
    //   .z  sorted position tid: camera | segment length << 22 at segment heads, -1 elsewhere
    //   .w  tile-local point tid: first observation | observation count << 16 (0 when tid >= points in tile)
    const int4* tmeta;            // [No]
    const int4* tile_info;        // [nTiles] {o0, n, first compact row, first local point}
    double* rows;                 // [nRows][ROW_DOUBLES]
    double* tile_intr;            // [nTiles][4] per-tile intrinsics-side partials (free intrinsics, one group)
    // by-camera view of the rows
    const int* cam_row_off;       // [Nc+1]
    const int* cam_rows;          // [nRows] compact row ids, ascending
    int nRows;
    // fused gather (BA_K5A_FUSE): rows delivered so far per camera; the CTA that delivers a camera's last row sums them all
    int* cam_cnt;                 // [Nc], zero between launches (the summing CTA resets its camera)
    const int* zero_cams;         // cameras without any row on this rank (nobody delivers them): CTA 0 stores their zeros
    int n_zero;
    int fuse;
};

// Sum of camera c's partial rows (static list, fixed order => bit-reproducible), by the whole CTA.  On return (after a CTA barrier)
// sw[0][0..5] holds the 6-vector.
// BA_GATHER_COOP (default): four adjacent lanes share one 64-byte row, lane k loading its k-th 16 bytes, so a warp instruction touches
// 8 rows = 8 lines.  The first version gave every lane a whole row (3 LDG.128 to 32 different lines per instruction): 3 L1
// wavefronts per row instead of 1, and ncu showed the kernel throttled by the load/store unit (lg_throttle 15, mio_throttle 12 per issue).
#ifndef BA_GATHER_COOP
#define BA_GATHER_COOP 1
#endif
#ifndef BA_GATHER_UNR
#define BA_GATHER_UNR 8        // rows in flight per lane
#endif
#ifndef BA_GATHER_MINB
#define BA_GATHER_MINB 12      // min CTAs/SM of the gather kernels: 12 = 40 registers (16 with BA_GATHER_UNR=4 keeps all 1779 Venice CTAs resident at once)
#endif
// WAIT: the kernel is a link of the PDL chain; the row list (static) is read first, griddepcontrol.wait comes right before the rows.
// Returns false (whole CTA, before any barrier) when *done says the PCG loop has already stopped.
template <bool WAIT = false>
__device__ __forceinline__ bool camera_row_sum(const RowSched& S, int c, double (*sw)[6], const int* done = nullptr) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int e = S.cam_row_off[c + 1];
#if BA_GATHER_COOP
    constexpr int NG = GATHER_WARPS * 8;          // 4-lane groups per CTA
    constexpr int UNR = BA_GATHER_UNR;
    const int sub = threadIdx.x & 3;
    const double2* R2 = reinterpret_cast<const double2*>(S.rows);
    double2 acc = make_double2(0.0, 0.0);
    int i = S.cam_row_off[c] + (threadIdx.x >> 2);
    int id[UNR];
    const bool full0 = i + (UNR - 1) * NG < e;
    if (full0) {
#pragma unroll
        for (int k = 0; k < UNR; ++k) id[k] = __ldg(S.cam_rows + i + k * NG);
    }
    if (WAIT) { pdl_wait(); if (done && *done) return false; }
    if (sub < 3) {
        for (bool first = true; i + (UNR - 1) * NG < e; i += UNR * NG, first = false) {     // UNR rows per group and trip, all in flight
            if (!first) {
#pragma unroll
                for (int k = 0; k < UNR; ++k) id[k] = __ldg(S.cam_rows + i + k * NG);
            }
            double2 v[UNR];
#pragma unroll
            for (int k = 0; k < UNR; ++k) v[k] = row_load(R2 + size_t(ROW_DOUBLES / 2) * id[k] + sub);
#pragma unroll
            for (int k = 1; k < UNR; ++k) { v[0].x += v[k].x; v[0].y += v[k].y; }
            acc.x += v[0].x; acc.y += v[0].y;
        }
        for (; i < e; i += NG) {
            const double2 v = row_load(R2 + size_t(ROW_DOUBLES / 2) * __ldg(S.cam_rows + i) + sub);
            acc.x += v.x; acc.y += v.y;
        }
    }
    // the 8 groups of a warp (lanes with equal sub), fixed tree
#pragma unroll
    for (int off = 4; off < 32; off <<= 1) { acc.x += __shfl_xor_sync(0xffffffffu, acc.x, off); acc.y += __shfl_xor_sync(0xffffffffu, acc.y, off); }
    __shared__ double swp[GATHER_WARPS][6];
    if (lane < 3) { swp[wid][2 * lane] = acc.x; swp[wid][2 * lane + 1] = acc.y; }
    __syncthreads();
    if (threadIdx.x < 6) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < GATHER_WARPS; ++w) v += swp[w][threadIdx.x];
        sw[0][threadIdx.x] = v;
    }
    __syncthreads();
#else
    if (WAIT) { pdl_wait(); if (done && *done) return false; }
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0, s5 = 0;
    // four independent rows per thread and trip: 4 index loads, then 12 row loads in flight
    constexpr int NT = GATHER_WARPS * 32;
    int i = S.cam_row_off[c] + threadIdx.x;
    for (; i + 3 * NT < e; i += 4 * NT) {
        const int i0 = S.cam_rows[i], i1 = S.cam_rows[i + NT], i2 = S.cam_rows[i + 2 * NT], i3 = S.cam_rows[i + 3 * NT];
        const double2* r0 = reinterpret_cast<const double2*>(S.rows + size_t(ROW_DOUBLES) * i0);
        const double2* r1 = reinterpret_cast<const double2*>(S.rows + size_t(ROW_DOUBLES) * i1);
        const double2* r2 = reinterpret_cast<const double2*>(S.rows + size_t(ROW_DOUBLES) * i2);
        const double2* r3 = reinterpret_cast<const double2*>(S.rows + size_t(ROW_DOUBLES) * i3);
        const double2 a0 = __ldcg(r0), b0 = __ldcg(r0 + 1), c0 = __ldcg(r0 + 2);
        const double2 a1 = __ldcg(r1), b1 = __ldcg(r1 + 1), c1 = __ldcg(r1 + 2);
        const double2 a2 = __ldcg(r2), b2 = __ldcg(r2 + 1), c2 = __ldcg(r2 + 2);
        const double2 a3 = __ldcg(r3), b3 = __ldcg(r3 + 1), c3 = __ldcg(r3 + 2);
        s0 += (a0.x + a1.x) + (a2.x + a3.x); s1 += (a0.y + a1.y) + (a2.y + a3.y);
        s2 += (b0.x + b1.x) + (b2.x + b3.x); s3 += (b0.y + b1.y) + (b2.y + b3.y);
        s4 += (c0.x + c1.x) + (c2.x + c3.x); s5 += (c0.y + c1.y) + (c2.y + c3.y);
    }
    for (; i < e; i += NT) {
        const double2* row = reinterpret_cast<const double2*>(S.rows + size_t(ROW_DOUBLES) * S.cam_rows[i]);
        const double2 a = __ldcg(row), bq = __ldcg(row + 1), cq = __ldcg(row + 2);
        s0 += a.x; s1 += a.y; s2 += bq.x; s3 += bq.y; s4 += cq.x; s5 += cq.y;
    }
    s0 = warp_sum(s0); s1 = warp_sum(s1); s2 = warp_sum(s2); s3 = warp_sum(s3); s4 = warp_sum(s4); s5 = warp_sum(s5);
    __shared__ double swp[GATHER_WARPS][6];
    if (lane == 0) { swp[wid][0] = s0; swp[wid][1] = s1; swp[wid][2] = s2; swp[wid][3] = s3; swp[wid][4] = s4; swp[wid][5] = s5; }
    __syncthreads();
    if (threadIdx.x < 6) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < GATHER_WARPS; ++w) v += swp[w][threadIdx.x];
        sw[0][threadIdx.x] = v;
    }
    __syncthreads();
#endif
    return true;
}

#ifndef BA_K5A_FUSE
#define BA_K5A_FUSE 0          // 1 / 2: compile the fused by-camera gather into K4/K5a (enabled at run time by RowSched::fuse). Measured on the
                               // Venice shape: 1 (fence + counter per row) K5a 205 -> 315 us against the 37 us gather it replaces; 2 = one fence per CTA
#endif
#ifndef BA_K5A_MINB
#define BA_K5A_MINB 6          // min CTAs/SM for ptxas: 6 = 80 registers, no spills (A/B knob: 7 and 8 cap the registers at 72 / 64)
#endif
#ifndef BA_K5A_XSMEM
#define BA_K5A_XSMEM 0        // 1: the search direction p[c] reaches the observations through shared memory, one fetch per distinct camera of the tile
#endif
#ifndef BA_K5A_OBS_HINV
#define BA_K5A_OBS_HINV 0      // 1: every observation thread sums its point's W and applies Hinv itself (one barrier less)
#endif
template <int MODE, bool FI, int PK>
__global__ void __launch_bounds__(TILE_OBS, BA_K5A_MINB)
k_schur_rows(View V, RowSched S, const double* __restrict__ xin, double* __restrict__ atomic_out, const int* __restrict__ done_flag) {
    __shared__ double sW[TILE_OBS][3];
    __shared__ double sV[TILE_OBS][3];
    __shared__ __align__(16) double sC[TILE_OBS][6];
    __shared__ double sred[32];
#if BA_K5A_FUSE
    __shared__ int s_ncomp, s_comp[TILE_OBS];
    __shared__ double s_sw[1][6];
    if (threadIdx.x == 0) s_ncomp = 0;      // several barriers before its first use
    // A camera's total is complete when its last row has been delivered; the deliverer's CTA then adds the rows in the static
    // order of the row list (so the sum does not depend on who does it) and stores the camera's 6-vector: the by-camera gather
    // (K5b, ~36 us as a kernel of its own on the Venice shape) runs in the shadow of the streaming sweep.
    auto deliver = [&](int cam) {
#if BA_K5A_FUSE == 1
        __threadfence();                                       // this thread's row first
#endif
        const int total = __ldg(S.cam_row_off + cam + 1) - __ldg(S.cam_row_off + cam);
        if (atomicAdd(S.cam_cnt + cam, 1) == total - 1) s_comp[atomicAdd(&s_ncomp, 1)] = cam;
    };
    auto finish = [&]() {                                      // whole CTA; between two CTA barriers
        __syncthreads();
        const int ncomp = s_ncomp;
        for (int k = 0; k < ncomp; ++k) {
            const int cam = s_comp[k];
            __threadfence();
            camera_row_sum(S, cam, s_sw);
            if (threadIdx.x < 6) __stcg(atomic_out + 6 * size_t(cam) + threadIdx.x, s_sw[0][threadIdx.x]);
            if (threadIdx.x == 0) S.cam_cnt[cam] = 0;
        }
        __syncthreads();
        if (threadIdx.x == 0) s_ncomp = 0;
    };
    auto zero_unseen = [&]() {                                 // after pdl_wait: the previous kernel of the chain wrote q over y
        if (!FI && S.fuse && blockIdx.x == 0)
            for (int i = threadIdx.x; i < 6 * S.n_zero; i += TILE_OBS) __stcg(atomic_out + 6 * size_t(S.zero_cams[i / 6]) + i % 6, 0.0);
    };
#endif
    pdl_trigger();
    const int t = blockIdx.x;
    const int4 ti = __ldg(S.tile_info + t);
    const int o0 = ti.x, n = ti.y, hbase = ti.z, pfirst = ti.w;
    const size_t ld = V.ldo;
    const int ioff = 6 * V.Nc;
    double yi[4] = {0, 0, 0, 0};

    if (n <= TILE_OBS) {
        // ---- every global load of this tile is issued here, before the first use (one latency, not five) ----
        const int o = o0 + threadIdx.x;
        const bool act = threadIdx.x < n;
        int4 tm = make_int4(0, 0, -1, 0);
        double2 jc[6], jp[3], ji[4], rr = make_double2(0.0, 0.0);
        if (act) {
            tm = ldg4s(S.tmeta + o);
            const double2* J = V.J2 + o;
            load_jc<PK>(V, J, jc);
            load_jp<PK>(V, J, jp);
            if (FI) load_ji<PK>(V, J, ji);
            if (MODE == SWEEP_RHS) rr = V.r2[o];
        }
        const int npt_obs = tm.w >> 16, pstart = tm.w & 0xffff;   // this thread as a point thread
        double2 h0 = make_double2(0, 0), h1 = h0, h2 = h0;
        if (MODE == SWEEP_MATVEC && npt_obs > 0 && !BA_K5A_OBS_HINV) {
            const double2* H = reinterpret_cast<const double2*>(V.Hinv + 6 * size_t(pfirst + threadIdx.x));
            h0 = __ldg(H); h1 = __ldg(H + 1); h2 = __ldg(H + 2);      // (H0,H1) (H2,H3) (H4,H5)
        }
        // everything above is static during a linear solve; xin and the done flag are the predecessor's (PCG tail) output
        pdl_wait();
        if (done_flag && *done_flag) return;
#if BA_K5A_FUSE
        zero_unseen();
#endif
        const int c = tm.x & 0xffffff, hp = int(unsigned(tm.x) >> 24);
        const int slot = tm.y & 0xff, n1 = (tm.y >> 8) & 0xff, hrank = (tm.y >> 16) & 0xff, lpt = (tm.y >> 24) & 0xff;
        const int pk = tm.z;
        const int g = (FI && act) ? V.cam_group[c] : 0;
        double2 u = make_double2(0.0, 0.0);
        double2 tt = make_double2(0.0, 0.0);
        if (MODE == SWEEP_MATVEC) {
#if BA_K5A_XSMEM
            // every distinct camera of the tile is fetched once, by its segment head (sorted position tid), into sC[tid]; the
            // observations read it back from there: 2.3x fewer divergent 16-byte gathers than one fetch per observation
            if (pk >= 0) {
                const double2* xc = reinterpret_cast<const double2*>(xin + 6 * (pk & 0x3fffff));
                double2* sx = reinterpret_cast<double2*>(sC[threadIdx.x]);
                sx[0] = __ldg(xc); sx[1] = __ldg(xc + 1); sx[2] = __ldg(xc + 2);
            }
            __syncthreads();
#endif
            if (act) {
#if BA_K5A_XSMEM
                const double2* sx = reinterpret_cast<const double2*>(sC[hp]);
                const double2 x0 = sx[0], x1 = sx[1], x2 = sx[2];
#else
                const double2* xc = reinterpret_cast<const double2*>(xin + 6 * c);
                const double2 x0 = __ldg(xc), x1 = __ldg(xc + 1), x2 = __ldg(xc + 2);
#endif
                u.x = jc[0].x * x0.x + jc[1].x * x0.y + jc[2].x * x1.x + jc[3].x * x1.y + jc[4].x * x2.x + jc[5].x * x2.y;
                u.y = jc[0].y * x0.x + jc[1].y * x0.y + jc[2].y * x1.x + jc[3].y * x1.y + jc[4].y * x2.x + jc[5].y * x2.y;
                if (FI) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) { const double xv = __ldg(xin + ioff + 4 * g + k); u.x += ji[k].x * xv; u.y += ji[k].y * xv; }
                }
#pragma unroll
                for (int a = 0; a < 3; ++a) sW[threadIdx.x][a] = jp[a].x * u.x + jp[a].y * u.y;
            }
#if BA_K5A_OBS_HINV
            reinterpret_cast<int*>(sV)[threadIdx.x] = tm.w;      // point thread's (first observation | count << 16), read back by its observations
            double2 g0 = make_double2(0, 0), g1 = g0, g2 = g0;
            if (act) {
                const double2* H = reinterpret_cast<const double2*>(V.Hinv + 6 * size_t(pfirst + lpt));
                g0 = __ldg(H); g1 = __ldg(H + 1); g2 = __ldg(H + 2);
            }
            __syncthreads();
            if (act) {
                const int pw = reinterpret_cast<const int*>(sV)[lpt];
                double W0 = 0, W1 = 0, W2 = 0;
                for (int i = pw & 0xffff, e = i + (pw >> 16); i < e; ++i) { W0 += sW[i][0]; W1 += sW[i][1]; W2 += sW[i][2]; }
                const double v0 = g0.x * W0 + g0.y * W1 + g1.x * W2, v1 = g0.y * W0 + g1.y * W1 + g2.x * W2, v2 = g1.x * W0 + g2.x * W1 + g2.y * W2;
                tt.x = u.x - (jp[0].x * v0 + jp[1].x * v1 + jp[2].x * v2);
                tt.y = u.y - (jp[0].y * v0 + jp[1].y * v1 + jp[2].y * v2);
            }
#else
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
#endif
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
            double2* dst = reinterpret_cast<double2*>(S.rows + size_t(ROW_DOUBLES) * (hbase + hrank));
            row_store(dst, a); row_store(dst + 1, bq); row_store(dst + 2, cq);
#if BA_K5A_FUSE == 1
            if (!FI && S.fuse) deliver(pk & 0x3fffff);
#endif
        }
#if BA_K5A_FUSE == 2
        if (!FI && S.fuse) {   // one fence for the whole CTA's rows (ordered before it by the barrier), then the counters
            __syncthreads();
            if (threadIdx.x == 0) __threadfence();
            __syncthreads();
            if (pk >= 0) deliver(pk & 0x3fffff);
        }
#endif
#if BA_K5A_FUSE
        if (!FI && S.fuse) finish();
#endif
    } else {
        // ---- one point with more than TILE_OBS observations (rare): two passes, every observation is its own row ----
        pdl_wait();
        if (done_flag && *done_flag) return;
#if BA_K5A_FUSE
        zero_unseen();
#endif
        const int o1 = o0 + n;
        const int p = pfirst;
        double v[3];
        if (MODE == SWEEP_MATVEC) {
            double W[3] = {0, 0, 0};
            for (int o = o0 + threadIdx.x; o < o1; o += TILE_OBS) {
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
#pragma unroll
            for (int a = 0; a < 3; ++a) { const double s = block_sum(W[a], sred); if (threadIdx.x == 0) sV[0][a] = s; }
            __syncthreads();
            const double* H = V.Hinv + 6 * size_t(p);
            const double w0 = sV[0][0], w1 = sV[0][1], w2 = sV[0][2];
            v[0] = H[0] * w0 + H[1] * w1 + H[2] * w2; v[1] = H[1] * w0 + H[3] * w1 + H[4] * w2; v[2] = H[2] * w0 + H[4] * w1 + H[5] * w2;
        } else { v[0] = V.vp[3 * size_t(p)]; v[1] = V.vp[3 * size_t(p) + 1]; v[2] = V.vp[3 * size_t(p) + 2]; }
        for (int base = o0; base < o1; base += TILE_OBS) {
          const int o = base + threadIdx.x;
          if (o < o1) {
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
#if BA_K5A_FUSE == 1
            if (!FI && S.fuse) deliver(c);
#endif
          }
#if BA_K5A_FUSE == 2
          if (!FI && S.fuse) {
              __syncthreads();
              if (threadIdx.x == 0) __threadfence();
              __syncthreads();
              if (o < o1) deliver(V.obs_cam[o]);
          }
#endif
#if BA_K5A_FUSE
          if (!FI && S.fuse) finish();
#endif
        }
    }
    if (FI && V.Ng == 1) {   // one shared intrinsics group: per-tile partial, summed in tile order by k_camera_gather
#pragma unroll
        for (int k = 0; k < 4; ++k) { const double s = block_sum(yi[k], sred); if (threadIdx.x == 0) S.tile_intr[4 * size_t(t) + k] = s; }
    }
}

// K5b — y[c] = sum of camera c's partial rows: one 4-warp CTA per camera, lanes stride the static row list, fixed combine
// order; with do_dot: q = y + Df2 p stored in y, one p.q partial per CTA, last CTA forms alpha = rho / p.q  (PcgState2).
// CTA `Nc` (the last) reduces the per-tile intrinsics partials instead (free intrinsics, single group).
__global__ void __launch_bounds__(GATHER_WARPS * 32, BA_GATHER_MINB)
k_camera_gather(int Nc, int Ng, int fi, int nTiles, RowSched S, const double* __restrict__ Df2, const double* __restrict__ p,
                double* __restrict__ y, double* __restrict__ partial, PcgState2* __restrict__ st, int do_dot) {
    __shared__ double sw[GATHER_WARPS][6];
    __shared__ bool last;
    pdl_trigger();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int ioff = 6 * Nc;
    double dot = 0.0;
    const int c = blockIdx.x;
    if (c >= Nc) { pdl_wait(); if (st && st->done) return; }   // PCG already stopped inside this batch of launches
    if (c < Nc) {
        if (!camera_row_sum<true>(S, c, sw, st ? &st->done : nullptr)) return;
        if (threadIdx.x < 6) {
            double v = sw[0][threadIdx.x];
            const int j = 6 * c + threadIdx.x;
            if (do_dot) { const double pj = p[j]; v += Df2[j] * pj; dot = pj * v; }
            y[j] = v;
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
        // fold the per-warp dots of this CTA into warp 0 lane 0 below
        __syncthreads();
    }
    if (!do_dot) return;
    // CTA partial of p.q: camera CTAs hold it in threads 0..5 (warp 0); the intrinsics CTA in lane 0 of each warp
    dot = warp_sum(dot);
    if (c >= Nc) { if (lane == 0) sw[wid][0] = dot; __syncthreads(); if (threadIdx.x == 0) { dot = 0.0; for (int w = 0; w < GATHER_WARPS; ++w) dot += sw[w][0]; } }
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = dot;
        __threadfence();
        const unsigned int tk = atomicAdd(&st->ticket_q, 1u);
        last = (tk == gridDim.x - 1);
    }
    __syncthreads();
    if (last) {   // fixed-order sum of the CTA partials by the whole last CTA (strided per thread, warp trees, then warp 0)
        __threadfence();
        double pq = 0.0;
        for (int i = threadIdx.x; i < int(gridDim.x); i += GATHER_WARPS * 32) pq += __ldcg(partial + i);
        pq = warp_sum(pq);
        __syncthreads();
        if (lane == 0) sw[wid][0] = pq;
        __syncthreads();
        pq = 0.0;
#pragma unroll
        for (int w = 0; w < GATHER_WARPS; ++w) pq += sw[w][0];
        if (wid == 0)
        if (lane == 0) {
            if (!(pq > 0.0) || !isfinite(pq)) { st->done = 3; st->alpha = 0.0; }
            else st->alpha = st->rho / pq;
            st->ticket_q = 0;
        }
    }
}

}  // namespace bak
