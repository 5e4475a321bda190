// ba_kernels.cuh — hand-written sm_100a kernels of the Levenberg-Marquardt inner loop (SURVEY.md §8 a-K).
//
// Everything is IEEE double (the reference's arithmetic type, BundleAdjuster.h:74-76); these are small-block streaming
// kernels bounded by HBM bandwidth, so no tensor cores.  Layout:
//   observations are sorted by point and cut into TILES of <= TILE_OBS observations holding whole points, one CTA per tile,
//   one thread per observation; a point's 3x3 work is done by one thread per point after a shared-memory exchange;
//   J is stored SoA as double2 column pairs  J2[k][o] = (dr0/dx_k, dr1/dx_k), k = 0..5 pose, 6..8 point, 9..12 intrinsics,
//   so every warp load is a fully coalesced 512-byte LDG.128 run;
//   per-camera data is a 160-byte record (R, t, aa and the Rodrigues-derivative coefficients) gathered through L1/L2.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bak {

#ifndef BA_TILE_OBS
#define BA_TILE_OBS 128   // 256 measured 3% slower on the Venice shape (K5a 222 vs 208 us: shorter barrier chains; K5b 54 vs 58 us: more rows), 64 slower again (K5b 66 us)
#endif
constexpr int TILE_OBS = BA_TILE_OBS;     // observations per tile == threads per CTA of the observation sweeps
constexpr int CAMREC = 20;        // doubles per camera record
constexpr int NJC = 6, NJP = 3, NJI = 4;

// camera record offsets
constexpr int CR_R = 0, CR_T = 9, CR_AA = 12, CR_A = 15, CR_B = 16, CR_C1 = 17, CR_C2 = 18, CR_C3 = 19;

struct RobustParams { int kind; double a; };

// Device-resident view of (this rank's shard of) the problem.  Point-indexed arrays are LOCAL to the shard.
struct View {
    int Nc, Ng, Np, No;          // cameras (global), groups (global), local points, local observations
    int ldo;                     // leading dimension (observations) of J2 / r2
    int nTiles;
    int fi;                      // 1: intrinsics blocks are free (J2 has 13 column pairs)
    int jpack;                   // 1: pinhole model, columns 3 and 4 share one stored column (see load_jc): J2 has 8 (12) column pairs
    int left_twist;              // 1: pose columns are d/d(omega, upsilon) of T <- exp(omega, upsilon) T (g2o policy); 0: d/d(angle-axis, t)
    int skip_jc;                 // K1 only: store neither the pose columns of J2 nor the by-observation records (nothing reads them when K2c/K4/K5/K7 run pose-factored)
    const int* tile_off;         // [nTiles+1] first observation of each tile
    const int* obs_cam;          // [No]
    const int* obs_pt;           // [No] local point id, non-decreasing
    const double2* obs_uv;       // [No]
    const int* pt_off;           // [Np+1]
    const int* cam_group;        // [Nc]
    const unsigned char* cam_free;  // [Nc] not constant and referenced
    const unsigned char* grp_free;  // [Ng]
    // camera-sorted view (for the by-camera segmented reductions)
    const int* perm_by_cam;      // [No] local observation ids, stable by camera
    const int* cam_off;          // [Nc+1]
    double2* J2;                 // [9 or 13][ldo]  ([8 or 12][ldo] with jpack)
    double2* r2;                 // [ldo] robustified residual
    double2* JcA;                // [No][8 or 12] by-observation records (jc0..jc5, r, [ji0..ji3], pad) for the by-camera gather of K2c
    int jca_stride;              // double2 per record: 8 (128 B), 12 with free intrinsics (192 B) — whole 32-byte sectors
    double* Hpp;                 // [Np][6] upper triangle xx xy xz yy yz zz of sum Jp^T Jp
    double* gp;                  // [Np][3] sum Jp^T r
    double* Hinv;                // [Np][6] (Hpp + Dp^2)^-1
    double* vp;                  // [Np][3] Hinv gp
    double* yp;                  // [Np][3] point part of the LM step
    double* sc_pt;               // [Np][3] Jacobi column scales
};

__device__ __forceinline__ double2 ldg2(const double2* p) { return __ldg(p); }
// Read-once streams (J columns, schedule records): do not allocate in L1, so the 85 KB search direction the same kernel gathers by
// camera stays L1-resident instead of being evicted by 900 MB of streaming data.  BA_J_NOALLOC=0 restores plain ld.global.nc.
#ifndef BA_J_NOALLOC
#define BA_J_NOALLOC 0
#endif
__device__ __forceinline__ double2 ldg2s(const double2* p) {
#if BA_J_NOALLOC
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
#else
    return __ldg(p);
#endif
}
__device__ __forceinline__ int4 ldg4s(const int4* p) {
#if BA_J_NOALLOC
    int4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
#else
    return __ldg(p);
#endif
}

// The Jacobian columns of observation o (J = V.J2 + o, column stride V.ldo).  Pinhole model (ReprojectCost, BundleAdjuster.h:55-62):
// u depends on P only through (Px, Pz) and v through (Py, Pz), so d(r0)/d(ty) = d(r1)/d(tx) = 0 exactly; with V.jpack the translation
// columns 3 = (a, 0) and 4 = (0, b) are stored as ONE column (a, b) and every later column moves up by one: 128 instead of 144 bytes per
// observation (192 / 208 with free intrinsics) in every sweep over J.
// PK: 1 packed, 0 unpacked (compile-time, the streaming kernels), -1 as V.jpack says (run-time, everything else).
template <int PK = -1>
__device__ __forceinline__ void load_jc(const View& V, const double2* __restrict__ J, double2 (&jc)[6]) {
    const size_t ld = V.ldo;
    const bool pk = PK < 0 ? V.jpack != 0 : PK != 0;
    jc[0] = ldg2s(J); jc[1] = ldg2s(J + ld); jc[2] = ldg2s(J + 2 * ld);
    const double2 d = ldg2s(J + 3 * ld);
    if (pk) { jc[3] = make_double2(d.x, 0.0); jc[4] = make_double2(0.0, d.y); jc[5] = ldg2s(J + 4 * ld); }
    else { jc[3] = d; jc[4] = ldg2s(J + 4 * ld); jc[5] = ldg2s(J + 5 * ld); }
}
template <int PK = -1>
__device__ __forceinline__ void load_jp(const View& V, const double2* __restrict__ J, double2 (&jp)[3]) {
    const int pk = PK < 0 ? V.jpack : PK;
    const double2* Jp = J + size_t(6 - pk) * V.ldo;
    jp[0] = ldg2s(Jp); jp[1] = ldg2s(Jp + V.ldo); jp[2] = ldg2s(Jp + 2 * size_t(V.ldo));
}
template <int PK = -1>
__device__ __forceinline__ void load_ji(const View& V, const double2* __restrict__ J, double2 (&ji)[4]) {
    const int pk = PK < 0 ? V.jpack : PK;
    const double2* Ji = J + size_t(9 - pk) * V.ldo;
#pragma unroll
    for (int k = 0; k < 4; ++k) ji[k] = ldg2s(Ji + size_t(k) * V.ldo);
}

__device__ __forceinline__ void red_add(double* addr, double v) {
    // fire-and-forget L2 reduction (RED.E.ADD.F64), no return value
    asm volatile("red.global.add.f64 [%0], %1;" ::"l"(addr), "d"(v) : "memory");
}

// per-thread asynchronous global->shared copies (LDGSTS); a thread that only reads back its OWN slots needs no barrier, just wait_all
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Programmatic dependent launch (ba_engine.cu: launch_chain).  No-ops in a kernel launched without the attribute.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide sum for blockDim.x <= 1024 (multiple of 32); result valid in thread 0. `scratch` holds >= 32 doubles.
__device__ __forceinline__ double block_sum(double v, double* scratch) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) scratch[wid] = v;
    __syncthreads();
    double s = 0.0;
    if (wid == 0) {
        s = lane < nw ? scratch[lane] : 0.0;
        s = warp_sum(s);
    }
    return s;
}

// Robust kernel on s = ||r||^2: returns rho(s), sets w = sqrt(rho'(s))   (ceres::HuberLoss / CauchyLoss)
__device__ __forceinline__ double robust(const RobustParams rp, double s, double& w) {
    const double b = rp.a * rp.a;
    if (rp.kind == 1) {
        if (s > b) {
            const double r = sqrt(s);
            const double rho1 = fmax(2.2250738585072014e-308, rp.a / r);
            w = sqrt(rho1);
            return 2.0 * rp.a * r - b;
        }
        w = 1.0;
        return s;
    } else if (rp.kind == 2) {
        const double sum = 1.0 + s / b;   // s * c with c = 1/b in Ceres; identical to rounding
        const double inv = 1.0 / sum;
        w = sqrt(fmax(2.2250738585072014e-308, inv));
        return b * log(sum);
    }
    w = 1.0;
    return s;
}

// ---------------------------------------------------------------------------------------------------------
// K0 — camera precompute: R(aa), and the coefficients of d(R(aa)X)/d(aa) in closed form:
//   R X = cos(th) X + a (aa x X) + b (aa.X) aa,  a = sin(th)/th, b = (1-cos th)/th^2
//   d(RX)/d(aa) = h aa^T - a [X]x + b ((aa.X) I + aa X^T),  h = c1 X + c2 (aa x X) + c3 (aa.X) aa
//   c1 = -a, c2 = (th cos th - sin th)/th^3, c3 = (th sin th - 2(1-cos th))/th^4   (power series below th^2 < 1e-2)
// Cameras on Ceres' first-order branch (AngleAxisRotatePoint) get R = I + [aa]x and d/daa = -[X]x exactly.
__global__ void k0_camera_precompute(int Nc, const double* __restrict__ pose, double* __restrict__ camrec, int taylor_rule) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= Nc) return;
    const double ax = pose[6 * c], ay = pose[6 * c + 1], az = pose[6 * c + 2];
    double* o = camrec + size_t(c) * CAMREC;
    const double th2 = ax * ax + ay * ay + az * az;
    bool taylor;
    if (taylor_rule == 0) taylor = !(th2 > 2.220446049250313e-16);
    else taylor = (ax == 0.0 && ay == 0.0 && az == 0.0);
    double a, b, c1, c2, c3, ct;
    if (taylor) {
        a = 1.0; b = 0.0; c1 = 0.0; c2 = 0.0; c3 = 0.0; ct = 1.0;
    } else {
        const double th = sqrt(th2);
        double sn, cs;
        sincos(th, &sn, &cs);
        ct = cs;
        if (th2 < 1e-2) {
            a = 1.0 + th2 * (-1.0 / 6 + th2 * (1.0 / 120 + th2 * (-1.0 / 5040 + th2 * (1.0 / 362880))));
            b = 0.5 + th2 * (-1.0 / 24 + th2 * (1.0 / 720 + th2 * (-1.0 / 40320 + th2 * (1.0 / 3628800))));
            c2 = -1.0 / 3 + th2 * (1.0 / 30 + th2 * (-1.0 / 840 + th2 * (1.0 / 45360 + th2 * (-1.0 / 3991680))));
            c3 = -1.0 / 12 + th2 * (1.0 / 180 + th2 * (-1.0 / 6720 + th2 * (1.0 / 453600 + th2 * (-1.0 / 47900160))));
        } else {
            a = sn / th;
            b = (1.0 - cs) / th2;
            c2 = (th * cs - sn) / (th2 * th);
            c3 = (th * sn - 2.0 * (1.0 - cs)) / (th2 * th2);
        }
        c1 = -a;
    }
    // R = ct I + a [aa]x + b aa aa^T
    o[CR_R + 0] = ct + b * ax * ax; o[CR_R + 1] = -a * az + b * ax * ay; o[CR_R + 2] = a * ay + b * ax * az;
    o[CR_R + 3] = a * az + b * ay * ax; o[CR_R + 4] = ct + b * ay * ay; o[CR_R + 5] = -a * ax + b * ay * az;
    o[CR_R + 6] = -a * ay + b * az * ax; o[CR_R + 7] = a * ax + b * az * ay; o[CR_R + 8] = ct + b * az * az;
    o[CR_T] = pose[6 * c + 3]; o[CR_T + 1] = pose[6 * c + 4]; o[CR_T + 2] = pose[6 * c + 5];
    o[CR_AA] = ax; o[CR_AA + 1] = ay; o[CR_AA + 2] = az;
    o[CR_A] = a; o[CR_B] = b; o[CR_C1] = c1; o[CR_C2] = c2; o[CR_C3] = c3;
}


// ---------------------------------------------------------------------------------------------------------
// A camera record into registers with 32-byte loads (LDG.E.256, new on sm_100; the record is 160 bytes and 32-byte aligned): the
// per-observation gathers are divergent (one record per lane, up to 32 different lines per warp instruction), so the number of load
// instructions, not the bytes, is what the L1 tag stage pays for — five instructions per record instead of ten with 16-byte loads.
__device__ __forceinline__ void ldg256(const double* p, double* d) {
    asm("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(d[0]), "=d"(d[1]), "=d"(d[2]), "=d"(d[3]) : "l"(p));
}
template <int NDOUBLE2>
__device__ __forceinline__ void load_camrec(const double* __restrict__ camrec, int c, double* __restrict__ dst) {
    static_assert(NDOUBLE2 % 2 == 0 && (CAMREC * 8) % 32 == 0, "whole 32-byte pieces of 32-byte aligned records");
    const double* src = camrec + size_t(c) * CAMREC;
#pragma unroll
    for (int k = 0; k < NDOUBLE2 / 2; ++k) ldg256(src + 4 * k, dst + 4 * k);
}

// Projection shared by K1 and K1': ReprojectCost::operator() (BundleAdjuster.h:38-68), trig-free thanks to K0.
struct Proj { double X[3]; double P[3]; double iz, xn, yn, r0, r1; };

// MODEL 0: ReprojectCost (BundleAdjuster.h:38-68), intrinsics (fx, fy, cx, cy), +z forward.
// MODEL 1: SnavelyReprojectionError (SnavelyReprojectionError.h:16-48), intrinsics (f, l1, l2), -z forward, radial distortion.
template <int MODEL>
__device__ __forceinline__ void project_obs(const double* __restrict__ cr, const double* __restrict__ X, const double* __restrict__ in,
                                            double2 uv, Proj& q) {
    q.X[0] = X[0]; q.X[1] = X[1]; q.X[2] = X[2];
    q.P[0] = cr[CR_R + 0] * X[0] + cr[CR_R + 1] * X[1] + cr[CR_R + 2] * X[2] + cr[CR_T + 0];
    q.P[1] = cr[CR_R + 3] * X[0] + cr[CR_R + 4] * X[1] + cr[CR_R + 5] * X[2] + cr[CR_T + 1];
    q.P[2] = cr[CR_R + 6] * X[0] + cr[CR_R + 7] * X[1] + cr[CR_R + 8] * X[2] + cr[CR_T + 2];
    q.iz = 1.0 / q.P[2];
    if (MODEL == 0) {
        q.xn = q.P[0] / q.P[2];
        q.yn = q.P[1] / q.P[2];
        q.r0 = in[0] * q.xn + in[2] - uv.x;   // u - observation.x   (:61,:64)
        q.r1 = in[1] * q.yn + in[3] - uv.y;   // v - observation.y   (:62,:65)
    } else {
        q.xn = -q.P[0] / q.P[2];              // xp, yp  (SnavelyReprojectionError.h:30-31)
        q.yn = -q.P[1] / q.P[2];
        const double r2 = q.xn * q.xn + q.yn * q.yn;
        const double dist = 1.0 + r2 * (in[1] + in[2] * r2);                 // :34-37
        q.r0 = in[0] * dist * q.xn - uv.x;                                   // :40-47
        q.r1 = in[0] * dist * q.yn - uv.y;
    }
}

// A = d(residual)/dP (2x3) and the intrinsics columns d(residual)/d(intrinsics) as (row0, row1) pairs (4 pairs; unused ones zero).
template <int MODEL>
__device__ __forceinline__ void model_jacobian(const Proj& q, const double* __restrict__ in, double A0[3], double A1[3], double2 Ji[4]) {
    if (MODEL == 0) {
        const double fx = in[0], fy = in[1];
        A0[0] = fx * q.iz; A0[1] = 0.0; A0[2] = -fx * q.xn * q.iz;
        A1[0] = 0.0; A1[1] = fy * q.iz; A1[2] = -fy * q.yn * q.iz;
        Ji[0] = make_double2(q.xn, 0.0); Ji[1] = make_double2(0.0, q.yn); Ji[2] = make_double2(1.0, 0.0); Ji[3] = make_double2(0.0, 1.0);
    } else {
        const double f = in[0], l1 = in[1], l2 = in[2];
        const double xp = q.xn, yp = q.yn, r2 = xp * xp + yp * yp;
        const double dist = 1.0 + r2 * (l1 + l2 * r2), dd = 2.0 * l1 + 4.0 * l2 * r2;
        // B = d(pred)/d(xp,yp) = f (dist I + dd (xp,yp)^T (xp,yp));  C = d(xp,yp)/dP = -[iz 0 -Px iz^2; 0 iz -Py iz^2] = [-iz 0 -xp iz; 0 -iz -yp iz]
        const double B00 = f * (dist + dd * xp * xp), B01 = f * dd * xp * yp, B11 = f * (dist + dd * yp * yp);
        A0[0] = -B00 * q.iz; A0[1] = -B01 * q.iz; A0[2] = -(B00 * xp + B01 * yp) * q.iz;
        A1[0] = -B01 * q.iz; A1[1] = -B11 * q.iz; A1[2] = -(B01 * xp + B11 * yp) * q.iz;
        Ji[0] = make_double2(dist * xp, dist * yp); Ji[1] = make_double2(f * r2 * xp, f * r2 * yp);
        Ji[2] = make_double2(f * r2 * r2 * xp, f * r2 * r2 * yp); Ji[3] = make_double2(0.0, 0.0);
    }
}

// K1 — residual + analytic Jacobian + robust weight + cost, fused with the by-point half of K2 (Hpp, g_p).
// One CTA per tile, one thread per observation; then one thread per point of the tile sums its observations'
// Jp^T Jp and Jp^T r from shared memory (observations are sorted by point, so a point's are contiguous).
// Algorithmic bytes: reads 24 No (uv, cam, pt ids) + 24 Np + 160 Nc; writes (144 [+64] + 16) No + 72 Np.
// Persistent: CTA b walks tiles b, b+G, ...; the NEXT tile's index data (tile offsets two tiles ahead; camera id, point id and
// pixel of this thread's observation one tile ahead) is loaded while the current tile is computed, so a tile's critical path
// starts at the camera-record / point gathers instead of three dependent DRAM latencies earlier.  One cost partial per CTA.
#ifndef BA_K1_JCA_SMEM
#define BA_K1_JCA_SMEM 1
#endif
#ifndef BA_K1_CAM_SMEM
#define BA_K1_CAM_SMEM 0   // needs BA_K1_JCA_SMEM (shares its staging area)
#endif
#ifndef K1_CTAS_PER_SM
#define K1_CTAS_PER_SM 4
#endif
// LEAN (engine: K2c/K4/K5/K7 all pose-factored): the pose columns and the by-observation records have no reader, so they are neither formed
// nor stored, and only R and t of the camera record are fetched (96 of its 160 bytes: three 32-byte loads per observation instead of five).
#ifndef K1_LEAN_CTAS_PER_SM
#define K1_LEAN_CTAS_PER_SM 6   // 4: 240 us, 5: 235 us, 6: 220 us on the Venice shape (112 -> 85 registers, no spills that matter)
#endif
template <int MODEL, bool LEAN = false>
__global__ void __launch_bounds__(TILE_OBS, LEAN ? K1_LEAN_CTAS_PER_SM : K1_CTAS_PER_SM)
k1_residual_jacobian(View V, const double* __restrict__ camrec, const double* __restrict__ intr, const double* __restrict__ pt,
                     RobustParams rp, double* __restrict__ cost_partial, int* __restrict__ bad_flag) {
    __shared__ double sJp[TILE_OBS][9];   // per observation: Jp (2x3) , r (2)  [+1 pad: stride 9 is conflict-free]
#if BA_K1_JCA_SMEM
    __shared__ __align__(16) double2 sA[TILE_OBS * 12];   // staging of the by-observation records (8 or 12 double2 each)
#endif
    __shared__ double sred[32];
    const int G = gridDim.x, nT = V.nTiles;
    int t = blockIdx.x;
    double cost = 0.0;
    int bad = 0;
    const int rot12 = (threadIdx.x >> 1) % 12;   // slot rotation of this thread's 12 x 16-byte staging area (sA)
    if (t >= nT) { if (threadIdx.x == 0) cost_partial[blockIdx.x] = 0.0; return; }
    int o0 = V.tile_off[t], o1 = V.tile_off[t + 1];
    // landing slots of the prefetch, double-buffered by tile parity: camera id, point id, pixel of this thread's observation
    __shared__ int sPc[2][TILE_OBS];
    __shared__ int sPp[2][TILE_OBS];
    __shared__ __align__(16) double2 sPuv[2][TILE_OBS];
    auto prefetch = [&](int buf, int a0, int a1) {
        if (a0 + int(threadIdx.x) < a1) {
            cp_async4(&sPc[buf][threadIdx.x], V.obs_cam + a0 + threadIdx.x);
            cp_async4(&sPp[buf][threadIdx.x], V.obs_pt + a0 + threadIdx.x);
            cp_async16(&sPuv[buf][threadIdx.x], V.obs_uv + a0 + threadIdx.x);
        }
        cp_async_commit();
    };
    int buf = 0;
    prefetch(0, o0, o1);
    int no0 = 0, no1 = 0;
    if (t + G < nT) { no0 = V.tile_off[t + G]; no1 = V.tile_off[t + G + 1]; }
    cp_async_wait_all();
    __syncthreads();
  for (;;) {
    // ---- this tile's index data is in sP*[buf] (visible to all: barrier at the end of the previous trip); start the copies for
    //      tile t+G into the other buffer and load the tile offsets of t+2G ----
    int nno0 = 0, nno1 = 0;
    if (t + G < nT) prefetch(buf ^ 1, no0, no1);
    if (t + 2 * G < nT) { nno0 = V.tile_off[t + 2 * G]; nno1 = V.tile_off[t + 2 * G + 1]; }
    const int pre_c = sPc[buf][threadIdx.x], pre_p = sPp[buf][threadIdx.x], pre_pf = sPp[buf][0];
    const double2 pre_uv = sPuv[buf][threadIdx.x];
    // Tiles normally hold <= TILE_OBS observations; a single point with more observations than that forms its own
    // tile and is walked in chunks (its Hpp/g_p are then accumulated across chunks by thread 0's running sums).
    double Hl[6] = {0, 0, 0, 0, 0, 0}, gl[3] = {0, 0, 0};   // long-point accumulators (thread 0 only)
    const bool longpt = (o1 - o0) > TILE_OBS;
    for (int base = o0; base < o1; base += TILE_OBS) {
        const int o = base + threadIdx.x;
        const bool act = o < o1;
        const bool first = base == o0;
        // by-point half, issued first so these dependent loads overlap the observation math: thread j takes the j-th point of the chunk
        const int lim = min(o1, base + TILE_OBS);
        const int pfirst = first ? pre_pf : V.obs_pt[base];
        const int plast = V.obs_pt[lim - 1];
        const int pj = pfirst + threadIdx.x;
        int pa = 0, pb = 0;
        if (pj <= plast) { pa = max(V.pt_off[pj], base); pb = min(V.pt_off[pj + 1], lim); }
#if BA_K1_CAM_SMEM
        {   // the chunk's camera records, fetched by consecutive lanes (10 lanes per 160-byte record: ~7 lines per warp instruction
            // instead of one line per lane) into the threads' staging areas; thread i then reads record i from its own area
            const int nq = (lim - base) * (CAMREC / 2);
            for (int q = threadIdx.x; q < nq; q += TILE_OBS) {
                const int i = q / (CAMREC / 2), k = q - i * (CAMREC / 2);
                const int cc = first ? sPc[buf][i] : V.obs_cam[base + i];
                int sl = k + (i >> 1) % 12; if (sl >= 12) sl -= 12;
                sA[i * 12 + sl] = __ldg(reinterpret_cast<const double2*>(camrec + size_t(cc) * CAMREC) + k);
            }
            __syncthreads();
        }
#endif
        if (act) {
            const int c = first ? pre_c : V.obs_cam[o], p = first ? pre_p : V.obs_pt[o];
            const int g = V.Ng == 1 ? 0 : V.cam_group[c];   // one shared Camera: no dependent group load
            double cr[CAMREC];
#if BA_K1_CAM_SMEM
#pragma unroll
            for (int k = 0; k < CAMREC / 2; ++k) { int sl = k + rot12; if (sl >= 12) sl -= 12; const double2 v = sA[threadIdx.x * 12 + sl]; cr[2 * k] = v.x; cr[2 * k + 1] = v.y; }
#else
            if (LEAN) load_camrec<6>(camrec, c, cr);
            else load_camrec<CAMREC / 2>(camrec, c, cr);
#endif
            const double* in = intr + 4 * g;
            Proj q;
            project_obs<MODEL>(cr, pt + 3 * size_t(p), in, first ? pre_uv : ldg2(V.obs_uv + o), q);
            const double s = q.r0 * q.r0 + q.r1 * q.r1;
            double w;
            const double rho = robust(rp, s, w);
            cost += 0.5 * rho;
            // A = d(residual)/dP; the pinhole model's structural zeros (A0[1], A1[0]) are literal constants and fold away
            double A0[3], A1[3];
            double2 Jin[4];
            model_jacobian<MODEL>(q, in, A0, A1, Jin);
            // J_X = A R
            double JX0[3], JX1[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                JX0[k] = A0[0] * cr[CR_R + k] + A0[1] * cr[CR_R + 3 + k] + A0[2] * cr[CR_R + 6 + k];
                JX1[k] = A1[0] * cr[CR_R + k] + A1[1] * cr[CR_R + 3 + k] + A1[2] * cr[CR_R + 6 + k];
            }
            double Ja0[3] = {0, 0, 0}, Ja1[3] = {0, 0, 0};
            if (!LEAN) {
            // J_aa = A G, row i: (A_i.h) aa - a (A_i x X) + b (d A_i + (A_i.aa) X)
            const double ax = cr[CR_AA], ay = cr[CR_AA + 1], az = cr[CR_AA + 2];
            const double X0 = q.X[0], X1 = q.X[1], X2 = q.X[2];
            const double d = ax * X0 + ay * X1 + az * X2;
            const double qx = ay * X2 - az * X1, qy = az * X0 - ax * X2, qz = ax * X1 - ay * X0;   // aa x X
            const double ca = cr[CR_A], cb = cr[CR_B], c1 = cr[CR_C1], c2 = cr[CR_C2], c3 = cr[CR_C3];
            const double hx = c1 * X0 + c2 * qx + c3 * d * ax, hy = c1 * X1 + c2 * qy + c3 * d * ay, hz = c1 * X2 + c2 * qz + c3 * d * az;
            if (V.left_twist) {   // g2o VertexSE3Expmap: dP/d(omega) = -[P]x  =>  row = P x A_i   (EdgeSE3ProjectXYZ::linearizeOplus)
                const double P0 = q.P[0], P1 = q.P[1], P2 = q.P[2];
                Ja0[0] = P1 * A0[2] - P2 * A0[1]; Ja0[1] = P2 * A0[0] - P0 * A0[2]; Ja0[2] = P0 * A0[1] - P1 * A0[0];
                Ja1[0] = P1 * A1[2] - P2 * A1[1]; Ja1[1] = P2 * A1[0] - P0 * A1[2]; Ja1[2] = P0 * A1[1] - P1 * A1[0];
            } else {
            {
                const double Ah = A0[0] * hx + A0[1] * hy + A0[2] * hz, Aa = A0[0] * ax + A0[1] * ay + A0[2] * az;
                const double cx0 = A0[1] * X2 - A0[2] * X1, cx1 = A0[2] * X0 - A0[0] * X2, cx2 = A0[0] * X1 - A0[1] * X0;  // A0 x X
                Ja0[0] = Ah * ax - ca * cx0 + cb * (d * A0[0] + Aa * X0);
                Ja0[1] = Ah * ay - ca * cx1 + cb * (d * A0[1] + Aa * X1);
                Ja0[2] = Ah * az - ca * cx2 + cb * (d * A0[2] + Aa * X2);
            }
            {
                const double Ah = A1[0] * hx + A1[1] * hy + A1[2] * hz, Aa = A1[0] * ax + A1[1] * ay + A1[2] * az;
                const double cx0 = A1[1] * X2 - A1[2] * X1, cx1 = A1[2] * X0 - A1[0] * X2, cx2 = A1[0] * X1 - A1[1] * X0;
                Ja1[0] = Ah * ax - ca * cx0 + cb * (d * A1[0] + Aa * X0);
                Ja1[1] = Ah * ay - ca * cx1 + cb * (d * A1[1] + Aa * X1);
                Ja1[2] = Ah * az - ca * cx2 + cb * (d * A1[2] + Aa * X2);
            }
            }
            }
            const double wc = V.cam_free[c] ? w : 0.0;   // constant pose: its Jacobian columns vanish (SetParameterBlockConstant, :105)
            double2* J = V.J2 + o;
            const size_t ld = V.ldo;
            const int pk = V.jpack;   // pinhole: A0[1] = A1[0] = 0, columns 3 and 4 share one stored column (load_jc)
            if (!LEAN && !V.skip_jc) {
                J[0 * ld] = make_double2(wc * Ja0[0], wc * Ja1[0]);
                J[1 * ld] = make_double2(wc * Ja0[1], wc * Ja1[1]);
                J[2 * ld] = make_double2(wc * Ja0[2], wc * Ja1[2]);
                if (pk) J[3 * ld] = make_double2(wc * A0[0], wc * A1[1]);
                else { J[3 * ld] = make_double2(wc * A0[0], wc * A1[0]); J[4 * ld] = make_double2(wc * A0[1], wc * A1[1]); }
                J[(5 - pk) * ld] = make_double2(wc * A0[2], wc * A1[2]);
            }
            const double p0[3] = {w * JX0[0], w * JX0[1], w * JX0[2]}, p1[3] = {w * JX1[0], w * JX1[1], w * JX1[2]};
            J[(6 - pk) * ld] = make_double2(p0[0], p1[0]);
            J[(7 - pk) * ld] = make_double2(p0[1], p1[1]);
            J[(8 - pk) * ld] = make_double2(p0[2], p1[2]);
            if (V.fi) {   // d(u,v)/d(fx,fy,cx,cy) = [x 0 1 0; 0 y 0 1]
                const double wi = V.grp_free[g] ? w : 0.0;
#pragma unroll
                for (int k = 0; k < 4; ++k) J[(9 - pk + k) * ld] = make_double2(wi * Jin[k].x, wi * Jin[k].y);
            }
            const double rr0 = w * q.r0, rr1 = w * q.r1;
            V.r2[o] = make_double2(rr0, rr1);
            if (!LEAN && !V.skip_jc) {   // by-observation record for K2c: whole sectors, so the by-camera gather fetches no byte it does not use (not written when K2c runs pose-factored)
#if BA_K1_JCA_SMEM
                // staged in shared memory (12 slots of 16 bytes per thread, chunk k in slot (k + rot) % 12: conflict-free for the 16-byte
                // accesses of 8 adjacent lanes) and written out after the barrier by consecutive lanes: whole lines per warp instruction.
                // Storing the records straight from the registers is 8 STG.128 to 32 different lines each = 256 L1 wavefronts per warp.
                double2* A = sA + threadIdx.x * 12;
                auto put = [&](int k, double2 v) { int sl = k + rot12; if (sl >= 12) sl -= 12; A[sl] = v; };
#else
                double2* A = V.JcA + size_t(o) * V.jca_stride;
                auto put = [&](int k, double2 v) { A[k] = v; };
#endif
                put(0, make_double2(wc * Ja0[0], wc * Ja1[0])); put(1, make_double2(wc * Ja0[1], wc * Ja1[1])); put(2, make_double2(wc * Ja0[2], wc * Ja1[2]));
                put(3, make_double2(wc * A0[0], wc * A1[0])); put(4, make_double2(wc * A0[1], wc * A1[1])); put(5, make_double2(wc * A0[2], wc * A1[2]));
                put(6, make_double2(rr0, rr1));
                if (V.fi) {
                    const double wi = V.grp_free[g] ? w : 0.0;
#pragma unroll
                    for (int k = 0; k < 4; ++k) put(7 + k, make_double2(wi * Jin[k].x, wi * Jin[k].y));
                    put(11, make_double2(0.0, 0.0));
                } else put(7, make_double2(0.0, 0.0));
            }
            double chk = rr0 + rr1 + rho;
#pragma unroll
            for (int k = 0; k < 3; ++k) chk += Ja0[k] + Ja1[k] + p0[k] + p1[k] + A0[k] + A1[k];
            if (!isfinite(chk)) bad = 1;
            double* sj = sJp[threadIdx.x];
            sj[0] = p0[0]; sj[1] = p0[1]; sj[2] = p0[2]; sj[3] = p1[0]; sj[4] = p1[1]; sj[5] = p1[2]; sj[6] = rr0; sj[7] = rr1;
        }
        __syncthreads();
#if BA_K1_JCA_SMEM
        if (!LEAN && !V.skip_jc) {   // write the chunk's records out: consecutive lanes take consecutive 16-byte pieces of the records
            const int st = V.jca_stride, nq = (lim - base) * st;
            double2* dst = V.JcA + size_t(base) * st;
            for (int q = threadIdx.x; q < nq; q += TILE_OBS) {
                const int i = st == 8 ? (q >> 3) : (q / 12), k = q - i * st;
                int sl = k + (i >> 1) % 12; if (sl >= 12) sl -= 12;
                dst[q] = sA[i * 12 + sl];
            }
        }
#endif
        // by-point half of K2
        const int p = pj;
        if (p <= plast) {
            const int a = pa, b = pb;
            double H[6] = {0, 0, 0, 0, 0, 0}, gg[3] = {0, 0, 0};
            for (int i = a; i < b; ++i) {
                const double* sj = sJp[i - base];
#pragma unroll
                for (int rrow = 0; rrow < 2; ++rrow) {
                    const double j0 = sj[3 * rrow], j1 = sj[3 * rrow + 1], j2 = sj[3 * rrow + 2], rv = sj[6 + rrow];
                    H[0] += j0 * j0; H[1] += j0 * j1; H[2] += j0 * j2; H[3] += j1 * j1; H[4] += j1 * j2; H[5] += j2 * j2;
                    gg[0] += j0 * rv; gg[1] += j1 * rv; gg[2] += j2 * rv;
                }
            }
            if (!longpt) {
#pragma unroll
                for (int k = 0; k < 6; ++k) V.Hpp[6 * size_t(p) + k] = H[k];
#pragma unroll
                for (int k = 0; k < 3; ++k) V.gp[3 * size_t(p) + k] = gg[k];
            } else {   // single long point: only thread 0 is active here
#pragma unroll
                for (int k = 0; k < 6; ++k) Hl[k] += H[k];
#pragma unroll
                for (int k = 0; k < 3; ++k) gl[k] += gg[k];
            }
        }
        __syncthreads();
    }
    if (longpt && threadIdx.x == 0) {
        const int p = V.obs_pt[o0];
#pragma unroll
        for (int k = 0; k < 6; ++k) V.Hpp[6 * size_t(p) + k] = Hl[k];
#pragma unroll
        for (int k = 0; k < 3; ++k) V.gp[3 * size_t(p) + k] = gl[k];
    }
    t += G;
    if (t >= nT) break;
    o0 = no0; o1 = no1; no0 = nno0; no1 = nno1;
    buf ^= 1;
    cp_async_wait_all();
    __syncthreads();   // every thread's copies for the new tile have landed and are visible (sPp[buf][0] is read by all)
  }
    if (bad) atomicOr(bad_flag, 1);
    const double cs = block_sum(cost, sred);
    if (threadIdx.x == 0) cost_partial[blockIdx.x] = cs;
}

// K1' — candidate cost: residual + rho only.  Reads 24 No + 24 Np + 160 Nc, writes one partial per CTA.
// Persistent grid-stride sweep (K1C_CTAS_PER_SM CTAs of 256 threads per SM): the observation's index data (camera, point, pixel) of the
// NEXT trip is loaded while the current trip's gathers (camera record, point) and arithmetic run, so a trip costs one gather latency, not
// two dependent ones.  (The first version was one short CTA per tile: 40 096 CTAs of three dependent loads and a block reduction each ran at
// 0.27 of the HBM roofline, bound by CTA turnover.)
constexpr int K1C_THREADS = 256;
constexpr int K1C_CTAS_PER_SM = 8;
template <int MODEL>
__global__ void __launch_bounds__(K1C_THREADS)
k1c_candidate_cost(View V, const double* __restrict__ camrec, const double* __restrict__ intr, const double* __restrict__ pt,
                   RobustParams rp, double* __restrict__ cost_partial, int* __restrict__ bad_flag) {
    __shared__ double sred[32];
    const int stride = gridDim.x * K1C_THREADS;
    double cost = 0.0;
    int bad = 0;
    int o = blockIdx.x * K1C_THREADS + threadIdx.x;
    int c = 0, p = 0;
    double2 uv = make_double2(0.0, 0.0);
    if (o < V.No) { c = __ldg(V.obs_cam + o); p = __ldg(V.obs_pt + o); uv = ldg2(V.obs_uv + o); }
    while (o < V.No) {
        const int on = o + stride;
        int cn = 0, pn = 0;
        double2 uvn = make_double2(0.0, 0.0);
        if (on < V.No) { cn = __ldg(V.obs_cam + on); pn = __ldg(V.obs_pt + on); uvn = ldg2(V.obs_uv + on); }
        Proj q;
        double cr[12];   // R, t
        load_camrec<6>(camrec, c, cr);
        project_obs<MODEL>(cr, pt + 3 * size_t(p), intr + (V.Ng == 1 ? 0 : 4 * V.cam_group[c]), uv, q);
        const double s = q.r0 * q.r0 + q.r1 * q.r1;
        double w;
        const double rho = robust(rp, s, w);
        if (!isfinite(rho)) bad = 1;
        cost += 0.5 * rho;
        o = on; c = cn; p = pn; uv = uvn;
    }
    if (bad) atomicOr(bad_flag, 1);
    const double cs = block_sum(cost, sred);
    if (threadIdx.x == 0) cost_partial[blockIdx.x] = cs;
}

// Deterministic final reduction of per-tile partials (fixed order for a fixed launch shape).
__global__ void k_reduce_partials(const double* __restrict__ partial, int n, double* __restrict__ out) {
    __shared__ double sred[32];
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += partial[i];
    s = block_sum(s, sred);
    if (threadIdx.x == 0) *out = s;
}


// ---------------------------------------------------------------------------------------------------------
// K4 / K5 / K7 — one sweep over the point-sorted Jacobian, three uses:
//   SWEEP_MATVEC : y += F^T (I - E (E^T E + Dp^2)^-1 E^T) F x      implicit Schur product, PCG inner loop (K5)
//   SWEEP_RHS    : b += F^T (r - E vp),  vp = (E^T E + Dp^2)^-1 E^T r   reduced right-hand side (K4)
//   SWEEP_BACKSUB: yp = vp - Hinv E^T F y;  partial of sum m.(r + m/2), m = -(F y + E yp)   (K7)
// F = [Jc | Ji] (pose and intrinsics columns), E = Jp.  J is read exactly once per sweep (144 [+64] B/observation);
// the camera-side scatter is a fire-and-forget L2 reduction (red.global.add.f64) per pose component.
// The damping term Df^2 x of the matvec is added by the PCG vector kernel, not here.
enum { SWEEP_MATVEC = 0, SWEEP_RHS = 1, SWEEP_BACKSUB = 2 };

template <int MODE, bool FI, int PK>
__global__ void __launch_bounds__(TILE_OBS)
k_schur_sweep(View V, const double* __restrict__ xin, double* __restrict__ yout, const int* __restrict__ done_flag,
              double* __restrict__ partial) {
    __shared__ double sW[TILE_OBS][3];
    __shared__ double sV[TILE_OBS][3];
    __shared__ double sred[32];
    __shared__ double sI[4];
    if (done_flag && *done_flag) return;   // PCG already converged: batched launches become no-ops
    const int t = blockIdx.x;
    const int o0 = V.tile_off[t], o1 = V.tile_off[t + 1];
    const int ioff = 6 * V.Nc;
    double acc = 0.0;                      // BACKSUB: model-cost partial
    double yi[4] = {0, 0, 0, 0};           // FI: intrinsics-side scatter, reduced per CTA when there is one group

    if (o1 - o0 <= TILE_OBS) {
        const int o = o0 + threadIdx.x;
        const bool act = o < o1;
        double2 jc[6], jp[3], ji[4];
        double2 u = make_double2(0.0, 0.0);
        int c = 0, p = 0, g = 0;
        if (act) {
            c = V.obs_cam[o]; p = V.obs_pt[o];
            const double2* J = V.J2 + o;
            load_jc<PK>(V, J, jc);
            load_jp<PK>(V, J, jp);
            if (FI) {
                g = V.cam_group[c];
                load_ji<PK>(V, J, ji);
            }
            if (MODE != SWEEP_RHS) {
                const double* xc = xin + 6 * c;      // (three 16-byte loads instead of these six measured 9 % slower in this kernel)
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
        if (MODE != SWEEP_RHS) {
            __syncthreads();
            const int pfirst = V.obs_pt[o0], plast = V.obs_pt[o1 - 1];
            const int pj = pfirst + threadIdx.x;
            if (pj <= plast) {
                double W[3] = {0, 0, 0};
                for (int i = V.pt_off[pj] - o0, e = V.pt_off[pj + 1] - o0; i < e; ++i) { W[0] += sW[i][0]; W[1] += sW[i][1]; W[2] += sW[i][2]; }
                const double* H = V.Hinv + 6 * size_t(pj);
                double v0 = H[0] * W[0] + H[1] * W[1] + H[2] * W[2];
                double v1 = H[1] * W[0] + H[3] * W[1] + H[4] * W[2];
                double v2 = H[2] * W[0] + H[4] * W[1] + H[5] * W[2];
                if (MODE == SWEEP_BACKSUB) {
                    v0 = V.vp[3 * size_t(pj)] - v0; v1 = V.vp[3 * size_t(pj) + 1] - v1; v2 = V.vp[3 * size_t(pj) + 2] - v2;
                    V.yp[3 * size_t(pj)] = v0; V.yp[3 * size_t(pj) + 1] = v1; V.yp[3 * size_t(pj) + 2] = v2;
                }
                sV[threadIdx.x][0] = v0; sV[threadIdx.x][1] = v1; sV[threadIdx.x][2] = v2;
            }
            __syncthreads();
            if (act) {
                const double* v = sV[p - pfirst];
                const double e0 = jp[0].x * v[0] + jp[1].x * v[1] + jp[2].x * v[2];
                const double e1 = jp[0].y * v[0] + jp[1].y * v[1] + jp[2].y * v[2];
                if (MODE == SWEEP_MATVEC) { tt.x = u.x - e0; tt.y = u.y - e1; }
                else {   // BACKSUB: m = -(F y + E yp)
                    const double m0 = -(u.x + e0), m1 = -(u.y + e1);
                    const double2 r = V.r2[o];
                    acc += m0 * (r.x + 0.5 * m0) + m1 * (r.y + 0.5 * m1);
                }
            }
        } else if (act) {
            const double* v = V.vp + 3 * size_t(p);
            const double2 r = V.r2[o];
            tt.x = r.x - (jp[0].x * v[0] + jp[1].x * v[1] + jp[2].x * v[2]);
            tt.y = r.y - (jp[0].y * v[0] + jp[1].y * v[1] + jp[2].y * v[2]);
        }
        if (MODE != SWEEP_BACKSUB && act) {
            double* yc = yout + 6 * c;
#pragma unroll
            for (int k = 0; k < 6; ++k) red_add(yc + k, jc[k].x * tt.x + jc[k].y * tt.y);
            if (FI) {
#pragma unroll
                for (int k = 0; k < 4; ++k) yi[k] = ji[k].x * tt.x + ji[k].y * tt.y;
                if (V.Ng > 1) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) red_add(yout + ioff + 4 * g + k, yi[k]);
                }
            }
        }
    } else {
        // ---- one point with more than TILE_OBS observations: two passes over its observations ----
        const int p = V.obs_pt[o0];
        double W[3] = {0, 0, 0};
        if (MODE != SWEEP_RHS) {
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
            if (threadIdx.x == 0) {
                const double* H = V.Hinv + 6 * size_t(p);
                const double w0 = sV[0][0], w1 = sV[0][1], w2 = sV[0][2];
                double v0 = H[0] * w0 + H[1] * w1 + H[2] * w2, v1 = H[1] * w0 + H[3] * w1 + H[4] * w2, v2 = H[2] * w0 + H[4] * w1 + H[5] * w2;
                if (MODE == SWEEP_BACKSUB) {
                    v0 = V.vp[3 * size_t(p)] - v0; v1 = V.vp[3 * size_t(p) + 1] - v1; v2 = V.vp[3 * size_t(p) + 2] - v2;
                    V.yp[3 * size_t(p)] = v0; V.yp[3 * size_t(p) + 1] = v1; V.yp[3 * size_t(p) + 2] = v2;
                }
                sV[1][0] = v0; sV[1][1] = v1; sV[1][2] = v2;
            }
            __syncthreads();
        }
        double v[3];
        if (MODE == SWEEP_RHS) { v[0] = V.vp[3 * size_t(p)]; v[1] = V.vp[3 * size_t(p) + 1]; v[2] = V.vp[3 * size_t(p) + 2]; }
        else { v[0] = sV[1][0]; v[1] = sV[1][1]; v[2] = sV[1][2]; }
        for (int o = o0 + threadIdx.x; o < o1; o += TILE_OBS) {
            const int c = V.obs_cam[o];
            const int g = FI ? V.cam_group[c] : 0;
            const double2* J = V.J2 + o;
            double2 jc[6], ji[4], jp[3];
            double2 u = make_double2(0.0, 0.0);
            load_jc<PK>(V, J, jc);
            if (FI) load_ji<PK>(V, J, ji);
            if (MODE != SWEEP_RHS) {
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
            else if (MODE == SWEEP_RHS) { const double2 r = V.r2[o]; tt.x = r.x - e0; tt.y = r.y - e1; }
            else {
                const double m0 = -(u.x + e0), m1 = -(u.y + e1);
                const double2 r = V.r2[o];
                acc += m0 * (r.x + 0.5 * m0) + m1 * (r.y + 0.5 * m1);
                continue;
            }
#pragma unroll
            for (int k = 0; k < 6; ++k) red_add(yout + 6 * c + k, jc[k].x * tt.x + jc[k].y * tt.y);
            if (FI) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const double contrib = ji[k].x * tt.x + ji[k].y * tt.y;
                    if (V.Ng > 1) red_add(yout + ioff + 4 * g + k, contrib); else yi[k] += contrib;
                }
            }
        }
    }
    if (FI && MODE != SWEEP_BACKSUB && V.Ng == 1) {   // one shared intrinsics group: reduce over the CTA, one RED per component
#pragma unroll
        for (int k = 0; k < 4; ++k) { const double s = block_sum(yi[k], sred); if (threadIdx.x == 0) sI[k] = s; }
        if (threadIdx.x == 0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) red_add(yout + ioff + k, sI[k]);
        }
    }
    if (MODE == SWEEP_BACKSUB) {
        const double s = block_sum(acc, sred);
        if (threadIdx.x == 0) partial[t] = s;
    }
}


// ---------------------------------------------------------------------------------------------------------
// K2c — by-camera half of K2: Hcc[c] = sum Jc^T Jc (6x6, stored full), gc[c] = sum Jc^T r, over the camera-sorted
// observation list (one CTA per camera, fixed summation order => deterministic).  With free intrinsics also the
// camera's share of its group's 4x4 block and gradient (ipart; k_group_blocks adds the shares in camera order).
// Gathers the by-observation records K1 wrote (JcA: Jc, r [, Ji] in 128 / 192 contiguous bytes = whole sectors) through
// perm_by_cam.  (v1 gathered the SoA columns: 7 separate 16-byte pieces per observation, 3.1x DRAM over-fetch.)  
#ifndef BA_K2C_COOP
#define BA_K2C_COOP 2   // 2: records staged through shared memory with cp.async (174 us on the Venice shape); 0: one record per lane (186 us); 1: 8 lanes per record + shuffles (296 us: 28 shuffles per record cost more than the 6 L1 wavefronts they save)
#endif
#if BA_K2C_COOP == 2
// Records staged through shared memory with cp.async (no staging registers): consecutive lanes fetch consecutive 16-byte pieces of the
// records (8 or 12 lanes per record: one L1 wavefront per record instead of 7 for one-record-per-lane) into rotated 16-byte slots
// (conflict-free for 8 adjacent lanes); every thread then reads ITS record back with 7 (11) LDS.128 and forms the 27 (41) products as the
// first version does.  Fixed intrinsics: two buffers, the next 128 records are in flight while this batch is multiplied.
#ifndef BA_K2C_MINB
#define BA_K2C_MINB 5
#endif
template <bool FI>
__global__ void __launch_bounds__(128, FI ? 4 : BA_K2C_MINB)
k2c_camera_blocks(View V, double* __restrict__ Hcc, double* __restrict__ gc, double* __restrict__ ipart) {
    constexpr int NV = FI ? (21 + 6 + 10 + 4) : (21 + 6);   // Hcc upper, gc, [Hii upper, gi]
    constexpr int ST = FI ? 12 : 8;                          // double2 per record
    constexpr int NBUF = FI ? 1 : 2;
    __shared__ double sm[4][NV];
    __shared__ __align__(16) double2 sR[NBUF][128 * ST];
    const int c = blockIdx.x;
    const int a0 = V.cam_off[c], a1 = V.cam_off[c + 1];
    auto rotof = [](int i) { return FI ? (i >> 1) % 12 : (i & 7); };
    const int rot = rotof(threadIdx.x);
    auto stage = [&](int buf, int ib) {
        const int nrec = min(128, a1 - ib);
#pragma unroll
        for (int m = 0; m < ST; ++m) {
            const int q = threadIdx.x + 128 * m;
            const int i = q / ST, k = q - i * ST;
            if (i < nrec) {
                const int o = __ldg(V.perm_by_cam + ib + i);
                int sl = k + rotof(i); if (sl >= ST) sl -= ST;
                cp_async16(&sR[buf][i * ST + sl], V.JcA + size_t(o) * ST + k);
            }
        }
        cp_async_commit();
    };
    double acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k] = 0.0;
    int buf = 0;
    if (a0 < a1) stage(0, a0);
    for (int ib = a0; ib < a1; ib += 128) {
        const int nrec = min(128, a1 - ib);
        cp_async_wait_all();
        __syncthreads();                                   // batch ib has landed; everybody is done with the other buffer
        if (NBUF == 2 && ib + 128 < a1) stage(buf ^ 1, ib + 128);
        if (threadIdx.x < nrec) {
            double2 j[FI ? 10 : 6], r;
            const double2* rec = sR[buf] + threadIdx.x * ST;
#pragma unroll
            for (int k = 0; k < 6; ++k) { int sl = k + rot; if (sl >= ST) sl -= ST; j[k] = rec[sl]; }
            { int sl = 6 + rot; if (sl >= ST) sl -= ST; r = rec[sl]; }
            if (FI) {
#pragma unroll
                for (int k = 0; k < 4; ++k) { int sl = 7 + k + rot; if (sl >= ST) sl -= ST; j[6 + k] = rec[sl]; }
            }
            int n = 0;
#pragma unroll
            for (int a = 0; a < 6; ++a)
#pragma unroll
                for (int b = a; b < 6; ++b) acc[n++] += j[a].x * j[b].x + j[a].y * j[b].y;
#pragma unroll
            for (int a = 0; a < 6; ++a) acc[n++] += j[a].x * r.x + j[a].y * r.y;
            if (FI) {
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = a; b < 4; ++b) acc[n++] += j[6 + a].x * j[6 + b].x + j[6 + a].y * j[6 + b].y;
#pragma unroll
                for (int a = 0; a < 4; ++a) acc[n++] += j[6 + a].x * r.x + j[6 + a].y * r.y;
            }
        }
        if (NBUF == 2) buf ^= 1;
        else { __syncthreads(); if (ib + 128 < a1) stage(0, ib + 128); }
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NV; ++k) { const double s = warp_sum(acc[k]); if (lane == 0) sm[wid][k] = s; }
    __syncthreads();
    if (threadIdx.x < NV) {
        const int k = threadIdx.x;
        const double s = sm[0][k] + sm[1][k] + sm[2][k] + sm[3][k];
        if (k < 21) {   // unpack upper-triangle index k -> (a,b)
            int a = 0, rem = k;
            while (rem >= 6 - a) { rem -= 6 - a; ++a; }
            const int b = a + rem;
            Hcc[36 * size_t(c) + 6 * a + b] = s;
            Hcc[36 * size_t(c) + 6 * b + a] = s;
        } else if (k < 27) {
            gc[6 * size_t(c) + (k - 21)] = s;
        } else if (FI) {
            ipart[14 * size_t(c) + (k - 27)] = s;
        }
    }
}
#elif BA_K2C_COOP
// 8 adjacent lanes (16 with free intrinsics) share one record: lane k loads its k-th 16 bytes, so a warp instruction reads 4 (2)
// whole records = 4 (3) lines, one L1 wavefront per record; the first version (below, #else) gave every lane a whole record: 7 LDG.128
// to 32 different lines each = 7 wavefronts per record, and the kernel sat at 0.55 of the HBM roofline waiting on the load/store unit.
// Lane a < 6 then forms row a of Jc^T Jc and (Jc^T r)[a] from the other lanes' pieces by width-8 shuffles (lanes 7..10: Ji).
template <bool FI>
__global__ void __launch_bounds__(128)
k2c_camera_blocks(View V, double* __restrict__ Hcc, double* __restrict__ gc, double* __restrict__ ipart) {
    constexpr int LPR = FI ? 16 : 8;        // lanes per record
    constexpr int NGRP = 128 / LPR;         // records per CTA and load instruction
    constexpr int UNR = 4;                  // records in flight per group
    __shared__ double sm[4][LPR][7];
    const int c = blockIdx.x;
    const int a0 = V.cam_off[c], a1 = V.cam_off[c + 1];
    const int sub = threadIdx.x & (LPR - 1), grp = threadIdx.x / LPR;
    const int sbase = (FI && sub >= 7) ? 7 : 0;
    const bool loads = sub < V.jca_stride;
    double acc[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) acc[k] = 0.0;
    for (int ib = a0; ib < a1; ib += UNR * NGRP) {   // the trip count must not depend on grp: the shuffles below are warp-wide
        const int i = ib + grp;
        int o[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) { const int idx = i + u * NGRP; o[u] = idx < a1 ? __ldg(V.perm_by_cam + idx) : -1; }
        double2 v[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) v[u] = (o[u] >= 0 && loads) ? ldg2(V.JcA + size_t(o[u]) * V.jca_stride + sub) : make_double2(0.0, 0.0);
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
#pragma unroll
            for (int bq = 0; bq < 6; ++bq) {
                const double jx = __shfl_sync(0xffffffffu, v[u].x, sbase + bq, LPR), jy = __shfl_sync(0xffffffffu, v[u].y, sbase + bq, LPR);
                acc[bq] += v[u].x * jx + v[u].y * jy;
            }
            const double rx = __shfl_sync(0xffffffffu, v[u].x, 6, LPR), ry = __shfl_sync(0xffffffffu, v[u].y, 6, LPR);
            acc[6] += v[u].x * rx + v[u].y * ry;
        }
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 7; ++k) {
#pragma unroll
        for (int off = LPR; off < 32; off <<= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], off);
    }
    if (lane < LPR) {
#pragma unroll
        for (int k = 0; k < 7; ++k) sm[wid][lane][k] = acc[k];
    }
    __syncthreads();
    if (threadIdx.x < LPR * 7) {
        const int a = threadIdx.x / 7, k = threadIdx.x % 7;
        const double s = (sm[0][a][k] + sm[1][a][k]) + (sm[2][a][k] + sm[3][a][k]);
        if (a < 6) {
            if (k < 6) Hcc[36 * size_t(c) + 6 * a + k] = s; else gc[6 * size_t(c) + a] = s;
        } else if (FI && a >= 7 && a < 11) {   // this camera's share of its group's 4x4 block (upper, 10) and gradient (4): k_group_blocks sums them
            const int ia = a - 7;
            if (k == 6) ipart[14 * size_t(c) + 10 + ia] = s;
            else if (k >= ia && k < 4) ipart[14 * size_t(c) + (ia * 4 - ia * (ia - 1) / 2) + (k - ia)] = s;
        }
    }
}
#else
template <bool FI>
__global__ void __launch_bounds__(128)
k2c_camera_blocks(View V, double* __restrict__ Hcc, double* __restrict__ gc, double* __restrict__ ipart) {
    constexpr int NV = FI ? (21 + 6 + 10 + 4) : (21 + 6);   // Hcc upper, gc, [Hii upper, gi]
    __shared__ double sm[4][NV];
    const int c = blockIdx.x;
    const int a0 = V.cam_off[c], a1 = V.cam_off[c + 1];
    double acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k] = 0.0;
    for (int i = a0 + threadIdx.x; i < a1; i += blockDim.x) {
        const int o = V.perm_by_cam[i];
        const double2* A = V.JcA + size_t(o) * V.jca_stride;   // one 128-byte (192-byte) record: whole sectors
        double2 j[FI ? 10 : 6];
#pragma unroll
        for (int k = 0; k < 6; ++k) j[k] = ldg2(A + k);
        const double2 r = ldg2(A + 6);
        if (FI) {
#pragma unroll
            for (int k = 0; k < 4; ++k) j[6 + k] = ldg2(A + 7 + k);
        }
        int n = 0;
#pragma unroll
        for (int a = 0; a < 6; ++a)
#pragma unroll
            for (int b = a; b < 6; ++b) acc[n++] += j[a].x * j[b].x + j[a].y * j[b].y;
#pragma unroll
        for (int a = 0; a < 6; ++a) acc[n++] += j[a].x * r.x + j[a].y * r.y;
        if (FI) {
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = a; b < 4; ++b) acc[n++] += j[6 + a].x * j[6 + b].x + j[6 + a].y * j[6 + b].y;
#pragma unroll
            for (int a = 0; a < 4; ++a) acc[n++] += j[6 + a].x * r.x + j[6 + a].y * r.y;
        }
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NV; ++k) { const double s = warp_sum(acc[k]); if (lane == 0) sm[wid][k] = s; }
    __syncthreads();
    if (threadIdx.x < NV) {
        const int k = threadIdx.x;
        const double s = sm[0][k] + sm[1][k] + sm[2][k] + sm[3][k];
        if (k < 21) {   // unpack upper-triangle index k -> (a,b)
            int a = 0, rem = k;
            while (rem >= 6 - a) { rem -= 6 - a; ++a; }
            const int b = a + rem;
            Hcc[36 * size_t(c) + 6 * a + b] = s;
            Hcc[36 * size_t(c) + 6 * b + a] = s;
        } else if (k < 27) {
            gc[6 * size_t(c) + (k - 21)] = s;
        } else if (FI) {
            ipart[14 * size_t(c) + (k - 27)] = s;   // this camera's share of its group's 4x4 block (upper, 10) and gradient (4): k_group_blocks sums them
        }
    }
}

#endif

// K3s — the pose blocks' Schur correction for preconditioner = BA_PRECOND_SCHUR_JACOBI (Ceres' SCHUR_JACOBI: the block diagonal of the
// reduced camera matrix S instead of that of B = F^T F + Dc^2):  corr[c] = sum over the points p camera c sees of W Hinv_p W^T with
// W = sum over c's observations of p of Jc^T Jp (6x3).  One CTA per camera over the camera-sorted list; that list is the stable sort of
// a point-sorted one, so repeated observations of a point by the same camera are adjacent and are folded by the first of them.
// Needs Hinv of THIS damping (runs after K3p); fixed-order reduction as K2c.  Bytes: (96 + 48 + 48 + 12) No + 288 Nc.
// EXPERIMENTAL: written in round 1 after the GPU budget was spent; compiled, not yet run (ba_create keeps it behind an opt-in).
__global__ void __launch_bounds__(128)
k_schur_diag(View V, double* __restrict__ corr) {
    __shared__ double sm[4][21];
    const int c = blockIdx.x;
    const int a0 = V.cam_off[c], a1 = V.cam_off[c + 1];
    double acc[21];
#pragma unroll
    for (int k = 0; k < 21; ++k) acc[k] = 0.0;
    for (int i = a0 + threadIdx.x; i < a1; i += blockDim.x) {
        const int o = V.perm_by_cam[i];
        const int p = V.obs_pt[o];
        if (i > a0 && V.obs_pt[V.perm_by_cam[i - 1]] == p) continue;      // a repeat: folded into the first observation of (c, p)
        double W[6][3];
#pragma unroll
        for (int a = 0; a < 6; ++a) { W[a][0] = 0.0; W[a][1] = 0.0; W[a][2] = 0.0; }
        for (int ii = i; ii < a1; ++ii) {
            const int oo = ii == i ? o : V.perm_by_cam[ii];
            if (ii != i && V.obs_pt[oo] != p) break;
            const double2* A = V.JcA + size_t(oo) * V.jca_stride;
            double2 jc[6], jp[3];
#pragma unroll
            for (int k = 0; k < 6; ++k) jc[k] = ldg2(A + k);
            load_jp(V, V.J2 + oo, jp);
#pragma unroll
            for (int a = 0; a < 6; ++a)
#pragma unroll
                for (int k = 0; k < 3; ++k) W[a][k] += jc[a].x * jp[k].x + jc[a].y * jp[k].y;
        }
        const double* h = V.Hinv + 6 * size_t(p);
        const double H[3][3] = {{h[0], h[1], h[2]}, {h[1], h[3], h[4]}, {h[2], h[4], h[5]}};
        double U[6][3];
#pragma unroll
        for (int a = 0; a < 6; ++a)
#pragma unroll
            for (int k = 0; k < 3; ++k) U[a][k] = W[a][0] * H[0][k] + W[a][1] * H[1][k] + W[a][2] * H[2][k];
        int n = 0;
#pragma unroll
        for (int a = 0; a < 6; ++a)
#pragma unroll
            for (int b = a; b < 6; ++b) acc[n++] += U[a][0] * W[b][0] + U[a][1] * W[b][1] + U[a][2] * W[b][2];
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 21; ++k) { const double s = warp_sum(acc[k]); if (lane == 0) sm[wid][k] = s; }
    __syncthreads();
    if (threadIdx.x < 21) {
        const int k = threadIdx.x;
        const double s = sm[0][k] + sm[1][k] + sm[2][k] + sm[3][k];
        int a = 0, rem = k;
        while (rem >= 6 - a) { rem -= 6 - a; ++a; }
        const int b = a + rem;
        corr[36 * size_t(c) + 6 * a + b] = s;
        corr[36 * size_t(c) + 6 * b + a] = s;
    }
}

// Free intrinsics: Hii / gi of every Camera group = sum of its cameras' shares, in camera order (one warp per group, fixed tree).
__global__ void __launch_bounds__(128)
k_group_blocks(int Ng, const int* __restrict__ grp_off, const int* __restrict__ grp_cams, const double* __restrict__ ipart,
               double* __restrict__ Hii, double* __restrict__ gi) {
    const int lane = threadIdx.x & 31;
    const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (g >= Ng) return;
    double acc[14];
#pragma unroll
    for (int k = 0; k < 14; ++k) acc[k] = 0.0;
    for (int i = grp_off[g] + lane; i < grp_off[g + 1]; i += 32) {
        const double* p = ipart + 14 * size_t(grp_cams[i]);
#pragma unroll
        for (int k = 0; k < 14; ++k) acc[k] += p[k];
    }
#pragma unroll
    for (int k = 0; k < 14; ++k) acc[k] = warp_sum(acc[k]);
    if (lane == 0) {
        int n = 0;
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = a; b < 4; ++b) { Hii[16 * size_t(g) + 4 * a + b] = acc[n]; Hii[16 * size_t(g) + 4 * b + a] = acc[n]; ++n; }
#pragma unroll
        for (int a = 0; a < 4; ++a) gi[4 * size_t(g) + a] = acc[10 + a];
    }
}

// Symmetric positive-definite n x n inverse by Cholesky on a thread-local array (n = 3, 4 or 6). false if not PD.
template <int N>
__device__ __forceinline__ bool spd_inverse(const double* A, double* Ainv) {
    double L[N * N];
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) {
            double s = A[i * N + j];
#pragma unroll
            for (int k = 0; k < j; ++k) s -= L[i * N + k] * L[j * N + k];
            if (i == j) { if (!(s > 0.0)) return false; L[i * N + i] = sqrt(s); }
            else L[i * N + j] = s / L[j * N + j];
        }
#pragma unroll
    for (int c = 0; c < N; ++c) {
        double y[N];
#pragma unroll
        for (int i = 0; i < N; ++i) { double s = (i == c) ? 1.0 : 0.0;
#pragma unroll
            for (int k = 0; k < i; ++k) s -= L[i * N + k] * y[k];
            y[i] = s / L[i * N + i]; }
#pragma unroll
        for (int i = N - 1; i >= 0; --i) { double s = y[i];
#pragma unroll
            for (int k = i + 1; k < N; ++k) s -= L[k * N + i] * Ainv[k * N + c];
            Ainv[i * N + c] = s / L[i * N + i]; }
    }
    return true;
}

struct LmParams { double radius, min_diag, max_diag; int policy; double lambda; };   // policy 1 (g2o): uniform damping lambda

// Damping in unscaled coordinates: Ceres solves in Jacobi-scaled space with D^2 = clamp(diag(Js^T Js))/radius;
// substituting Js = J diag(s) gives the identical step from (J^T J + D'^2) y = J^T r with
//   D'^2_j = clamp(s_j^2 H_jj, min, max) / (s_j^2 radius).
__device__ __forceinline__ double damping(double Hjj, double s, const LmParams lm) {
    if (lm.policy == 1) return lm.lambda;
    const double s2 = s * s;
    return fmin(fmax(s2 * Hjj, lm.min_diag), lm.max_diag) / (s2 * lm.radius);
}

// K3p — per point: Hinv = (Hpp + Dp'^2)^-1 (closed form through Cholesky), vp = Hinv gp.  168 Np bytes.
__global__ void k3_point_blocks(View V, LmParams lm, int* __restrict__ bad_flag) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= V.Np) return;
    const double2* Hv = reinterpret_cast<const double2*>(V.Hpp + 6 * size_t(p));   // 48-byte rows: three 16-byte loads
    const double2 ha = Hv[0], hb = Hv[1], hc = Hv[2];
    const double H[6] = {ha.x, ha.y, hb.x, hb.y, hc.x, hc.y};
    const double* s = V.sc_pt + 3 * size_t(p);
    double A[9];
    A[0] = H[0] + damping(H[0], s[0], lm); A[1] = H[1]; A[2] = H[2];
    A[3] = H[1]; A[4] = H[3] + damping(H[3], s[1], lm); A[5] = H[4];
    A[6] = H[2]; A[7] = H[4]; A[8] = H[5] + damping(H[5], s[2], lm);
    double I[9];
    if (!spd_inverse<3>(A, I)) { atomicOr(bad_flag, 2); for (int k = 0; k < 9; ++k) I[k] = 0.0; }
    double2* o = reinterpret_cast<double2*>(V.Hinv + 6 * size_t(p));
    o[0] = make_double2(I[0], I[1]); o[1] = make_double2(I[2], I[4]); o[2] = make_double2(I[5], I[8]);
    const double* g = V.gp + 3 * size_t(p);
    V.vp[3 * size_t(p)] = I[0] * g[0] + I[1] * g[1] + I[2] * g[2];
    V.vp[3 * size_t(p) + 1] = I[3] * g[0] + I[4] * g[1] + I[5] * g[2];
    V.vp[3 * size_t(p) + 2] = I[6] * g[0] + I[7] * g[1] + I[8] * g[2];
}

// K3c — per pose (and per intrinsics group): damping diagonal Df2 and block-Jacobi preconditioner
// Minv = (Hcc + Dc'^2)^-1.  Block b < Nc is a pose (6x6 at Minv[36 b]); block Nc+g is a group (4x4 at Minv[36 Nc + 16 g]).
__global__ void k3_camera_blocks(int Nc, int Ng, int fi, const double* __restrict__ Hcc, const double* __restrict__ Hii,
                                 const double* __restrict__ sc_f, LmParams lm, double* __restrict__ Df2, double* __restrict__ Minv,
                                 int* __restrict__ bad_flag) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < Nc) {
        double A[36], I[36];
        for (int k = 0; k < 36; ++k) A[k] = Hcc[36 * size_t(b) + k];
        for (int k = 0; k < 6; ++k) { const double d = damping(A[7 * k], sc_f[6 * b + k], lm); Df2[6 * b + k] = d; A[7 * k] += d; }
        if (!Minv) return;   // direct solve: only the damping is needed
        if (!spd_inverse<6>(A, I)) { atomicOr(bad_flag, 4); for (int k = 0; k < 36; ++k) I[k] = 0.0; }
        for (int k = 0; k < 36; ++k) Minv[36 * size_t(b) + k] = I[k];
    } else if (fi && b < Nc + Ng) {
        const int g = b - Nc;
        double A[16], I[16];
        for (int k = 0; k < 16; ++k) A[k] = Hii[16 * size_t(g) + k];
        for (int k = 0; k < 4; ++k) { const double d = damping(A[5 * k], sc_f[6 * Nc + 4 * g + k], lm); Df2[6 * Nc + 4 * g + k] = d; A[5 * k] += d; }
        if (!Minv) return;
        if (!spd_inverse<4>(A, I)) { atomicOr(bad_flag, 4); for (int k = 0; k < 16; ++k) I[k] = 0.0; }
        for (int k = 0; k < 16; ++k) Minv[36 * size_t(Nc) + 16 * size_t(g) + k] = I[k];
    }
}

// Jacobi column scales s_j = 1/(1 + sqrt(H_jj)) from the iteration-0 Hessian diagonals (frozen afterwards).
// Schur-Jacobi pose blocks: overwrites k3_camera_blocks' Minv[c] by (Hcc[c] + Dc^2 - corr[c])^-1; if rounding made a block indefinite
// the plain block (already in Minv) stays.  Runs after k3_camera_blocks (Df2) and k_schur_diag (corr).  EXPERIMENTAL, see k_schur_diag.
__global__ void k3_schur_jacobi_blocks(int Nc, const double* __restrict__ Hcc, const double* __restrict__ Df2, const double* __restrict__ corr,
                                       double* __restrict__ Minv) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= Nc) return;
    double A[36], I[36];
    for (int k = 0; k < 36; ++k) A[k] = Hcc[36 * size_t(b) + k] - corr[36 * size_t(b) + k];
    for (int k = 0; k < 6; ++k) A[7 * k] += Df2[6 * b + k];
    if (spd_inverse<6>(A, I)) for (int k = 0; k < 36; ++k) Minv[36 * size_t(b) + k] = I[k];
}

__global__ void k_scales_points(View V, int enabled) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 3 * V.Np) return;
    const int p = i / 3, a = i % 3;
    const int d = (a == 0) ? 0 : (a == 1 ? 3 : 5);
    V.sc_pt[i] = enabled ? 1.0 / (1.0 + sqrt(V.Hpp[6 * size_t(p) + d])) : 1.0;
}
__global__ void k_scales_cameras(int Nc, int Ng, int fi, const double* __restrict__ Hcc, const double* __restrict__ Hii,
                                 double* __restrict__ sc_f, int enabled) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 6 * Nc) sc_f[i] = enabled ? 1.0 / (1.0 + sqrt(Hcc[36 * size_t(i / 6) + 7 * (i % 6)])) : 1.0;
    else if (fi && i < 6 * Nc + 4 * Ng) { const int j = i - 6 * Nc; sc_f[i] = enabled ? 1.0 / (1.0 + sqrt(Hii[16 * size_t(j / 4) + 5 * (j % 4)])) : 1.0; }
}


// ---------------------------------------------------------------------------------------------------------
// PCG on the reduced camera system (block-Jacobi preconditioner, implicit Schur product by the K5 sweep).
// All vectors have nf = 6 Nc (+ 4 Ng) entries (85 KB on Venice) and are replicated on every rank; the kernels are in
// ba_scatter.cuh.  No scalar travels to the host inside the loop: the stop test lives in PcgState2.done.
struct PcgState2 {
    double rho, rho0, rel, Q0, alpha, beta;
    int k, done;                  // done: 0 running, 1 converged, 2 iteration cap, 3 breakdown
    unsigned int ticket_q, ticket_r;
};
struct PcgParams { int max_iterations; int forcing; double rel_tolerance; double eta; };

struct Scalars {   // everything the host reads back, one D2H copy
    double cost, ccost, mcc, step2_pt, xn2_pt, step2_f, xn2_f, gmax_pt, gmax_f;
    PcgState2 pcg;
    int bad_eval, bad_cand, bad_lin, pad;
    double maxdiag_pt, maxdiag_f, scale_pt, scale_f;   // g2o policy: max diagonal of H (lambda init), dx.(lambda dx + b)
};

__device__ __forceinline__ double apply_minv(int j, int Nc, const double* __restrict__ Minv, const double* __restrict__ r) {
    if (j < 6 * Nc) {
        const int c = j / 6, a = j - 6 * c;
        const double* M = Minv + 36 * size_t(c) + 6 * a;
        const double* rc = r + 6 * c;
        return M[0] * rc[0] + M[1] * rc[1] + M[2] * rc[2] + M[3] * rc[3] + M[4] * rc[4] + M[5] * rc[5];
    }
    const int jj = j - 6 * Nc, g = jj / 4, a = jj - 4 * g;
    const double* M = Minv + 36 * size_t(Nc) + 16 * size_t(g) + 4 * a;
    const double* rg = r + 6 * Nc + 4 * g;
    return M[0] * rg[0] + M[1] * rg[1] + M[2] * rg[2] + M[3] * rg[3];
}

// ---------------------------------------------------------------------------------------------------------
// K8 — candidate x+ = x - y (the step is delta = -y), with the norms the minimizer needs:
//   step2 = sum (x - x+)^2 formed from the rounded difference exactly as Ceres does, xn2 = sum (x+)^2, free scalars only.
__global__ void __launch_bounds__(256)
k8_candidate_points(View V, const double* __restrict__ pt, double* __restrict__ cpt, double* __restrict__ step2_partial,
                    double* __restrict__ xn2_partial, double* __restrict__ scale_partial, double lambda) {
    __shared__ double sred[32];
    double s2 = 0.0, x2 = 0.0, sc = 0.0;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < V.Np; p += gridDim.x * blockDim.x) {   // one point per thread and trip
        const bool fr = V.pt_off[p + 1] > V.pt_off[p];
        const size_t i = 3 * size_t(p);
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const double xo = pt[i + a], y = V.yp[i + a];
            const double xc = xo - y;
            cpt[i + a] = xc;
            if (fr) { const double d = xo - xc; s2 += d * d; x2 += xc * xc; sc += y * (lambda * y + V.gp[i + a]); }
        }
    }
    s2 = block_sum(s2, sred);
    if (threadIdx.x == 0) step2_partial[blockIdx.x] = s2;
    x2 = block_sum(x2, sred);
    if (threadIdx.x == 0) xn2_partial[blockIdx.x] = x2;
    sc = block_sum(sc, sred);
    if (threadIdx.x == 0) scale_partial[blockIdx.x] = sc;
}

// SO(3) exponential / logarithm and the left SE(3) update of g2o's VertexSE3Expmap::oplus: T <- exp(omega, upsilon) T
__device__ __forceinline__ void so3_exp_dev(const double w[3], double R[9]) {
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = sqrt(th2);
    double a, b;
    if (th < 1e-5) { a = 1.0 - th2 / 6.0; b = 0.5 - th2 / 24.0; } else { a = sin(th) / th; b = (1.0 - cos(th)) / th2; }
    R[0] = 1.0 - b * (w[1] * w[1] + w[2] * w[2]); R[1] = -a * w[2] + b * w[0] * w[1]; R[2] = a * w[1] + b * w[0] * w[2];
    R[3] = a * w[2] + b * w[0] * w[1]; R[4] = 1.0 - b * (w[0] * w[0] + w[2] * w[2]); R[5] = -a * w[0] + b * w[1] * w[2];
    R[6] = -a * w[1] + b * w[0] * w[2]; R[7] = a * w[0] + b * w[1] * w[2]; R[8] = 1.0 - b * (w[0] * w[0] + w[1] * w[1]);
}
__device__ __forceinline__ void so3_log_dev(const double R[9], double w[3]) {
    double q[4];
    const double tr = R[0] + R[4] + R[8];
    if (tr > 0.0) { double t = sqrt(tr + 1.0); q[0] = 0.5 * t; t = 0.5 / t; q[1] = (R[7] - R[5]) * t; q[2] = (R[2] - R[6]) * t; q[3] = (R[3] - R[1]) * t; }
    else {
        int i = 0; if (R[4] > R[0]) i = 1; if (R[8] > R[4 * i]) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        double t = sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
        q[1 + i] = 0.5 * t; t = 0.5 / t;
        q[0] = (R[3 * k + j] - R[3 * j + k]) * t; q[1 + j] = (R[3 * j + i] + R[3 * i + j]) * t; q[1 + k] = (R[3 * k + i] + R[3 * i + k]) * t;
    }
    if (q[0] < 0.0) { q[0] = -q[0]; q[1] = -q[1]; q[2] = -q[2]; q[3] = -q[3]; }
    const double n = sqrt(q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    if (n == 0.0) { w[0] = w[1] = w[2] = 0.0; return; }
    const double ang = 2.0 * atan2(n, q[0]);
    w[0] = ang * q[1] / n; w[1] = ang * q[2] / n; w[2] = ang * q[3] / n;
}
__device__ __forceinline__ void se3_left_update_dev(const double* pose, const double w[3], const double u[3], double* out) {
    double Rd[9], R[9], Rn[9];
    const double aa[3] = {pose[0], pose[1], pose[2]};
    so3_exp_dev(w, Rd); so3_exp_dev(aa, R);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) Rn[i * 3 + j] = Rd[i * 3] * R[j] + Rd[i * 3 + 1] * R[3 + j] + Rd[i * 3 + 2] * R[6 + j];
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = sqrt(th2);
    double b, c;
    if (th < 1e-5) { b = 0.5 - th2 / 24.0; c = 1.0 / 6.0 - th2 / 120.0; } else { b = (1.0 - cos(th)) / th2; c = (th - sin(th)) / (th2 * th); }
    const double wu[3] = {w[1] * u[2] - w[2] * u[1], w[2] * u[0] - w[0] * u[2], w[0] * u[1] - w[1] * u[0]};
    const double wwu[3] = {w[1] * wu[2] - w[2] * wu[1], w[2] * wu[0] - w[0] * wu[2], w[0] * wu[1] - w[1] * wu[0]};
    double lg[3];
    so3_log_dev(Rn, lg);
    out[0] = lg[0]; out[1] = lg[1]; out[2] = lg[2];
#pragma unroll
    for (int i = 0; i < 3; ++i) out[3 + i] = Rd[i * 3] * pose[3] + Rd[i * 3 + 1] * pose[4] + Rd[i * 3 + 2] * pose[5] + u[i] + b * wu[i] + c * wwu[i];
}

__global__ void __launch_bounds__(1024)
k8_candidate_cameras(int Nc, int Ng, int fi, const unsigned char* __restrict__ cam_free, const unsigned char* __restrict__ grp_free,
                     const double* __restrict__ pose, const double* __restrict__ intr, const double* __restrict__ yf,
                     double* __restrict__ cpose, double* __restrict__ cintr, const double* __restrict__ step2_partial,
                     const double* __restrict__ xn2_partial, int npartial, Scalars* __restrict__ sc,
                     const double* __restrict__ scale_partial, const double* __restrict__ gc, double lambda, int left_twist) {
    __shared__ double sred[32];
    double s2 = 0.0, x2 = 0.0, ps = 0.0, px = 0.0, sf = 0.0, sp = 0.0;
    if (left_twist) {   // g2o policy: T <- exp(-y_omega, -y_upsilon) T per free pose; dx.(lambda dx + b) with dx = -y, b = -g
        for (int c = threadIdx.x; c < Nc; c += blockDim.x) {
            if (cam_free[c]) {
                const double w[3] = {-yf[6 * c], -yf[6 * c + 1], -yf[6 * c + 2]}, u[3] = {-yf[6 * c + 3], -yf[6 * c + 4], -yf[6 * c + 5]};
                se3_left_update_dev(pose + 6 * c, w, u, cpose + 6 * c);
#pragma unroll
                for (int k = 0; k < 6; ++k) { const double y = yf[6 * c + k]; sf += y * (lambda * y + gc[6 * c + k]); const double d = pose[6 * c + k] - cpose[6 * c + k]; s2 += d * d; x2 += cpose[6 * c + k] * cpose[6 * c + k]; }
            } else {
#pragma unroll
                for (int k = 0; k < 6; ++k) cpose[6 * c + k] = pose[6 * c + k];
            }
        }
    } else
    for (int j = threadIdx.x; j < 6 * Nc; j += blockDim.x) {
        const double xo = pose[j];
        if (cam_free[j / 6]) { const double xc = xo - yf[j]; cpose[j] = xc; const double d = xo - xc; s2 += d * d; x2 += xc * xc; }
        else cpose[j] = xo;
    }
    for (int j = threadIdx.x; j < 4 * Ng; j += blockDim.x) {
        const double xo = intr[j];
        if (fi && grp_free[j / 4]) { const double xc = xo - yf[6 * Nc + j]; cintr[j] = xc; const double d = xo - xc; s2 += d * d; x2 += xc * xc; }
        else cintr[j] = xo;
    }
    for (int j = threadIdx.x; j < npartial; j += blockDim.x) { ps += step2_partial[j]; px += xn2_partial[j]; sp += scale_partial[j]; }   // point-side partials (fixed order)
    s2 = block_sum(s2, sred);
    if (threadIdx.x == 0) sc->step2_f = s2;
    x2 = block_sum(x2, sred);
    if (threadIdx.x == 0) sc->xn2_f = x2;
    ps = block_sum(ps, sred);
    if (threadIdx.x == 0) sc->step2_pt = ps;
    px = block_sum(px, sred);
    if (threadIdx.x == 0) sc->xn2_pt = px;
    sf = block_sum(sf, sred);
    if (threadIdx.x == 0) sc->scale_f = sf;
    sp = block_sum(sp, sred);
    if (threadIdx.x == 0) sc->scale_pt = sp;
}

// g2o computeLambdaInit: max diagonal entry of H over point blocks (per-block partial) and pose blocks (final kernel).
__global__ void __launch_bounds__(256)
k_maxdiag_points(View V, double* __restrict__ max_partial) {
    __shared__ double smax[32];
    double m = 0.0;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < V.Np; p += gridDim.x * blockDim.x) {
        const double* H = V.Hpp + 6 * size_t(p);
        m = fmax(m, fmax(fabs(H[0]), fmax(fabs(H[3]), fabs(H[5]))));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) { double r = 0.0; for (int w = 0; w < (blockDim.x >> 5); ++w) r = fmax(r, smax[w]); max_partial[blockIdx.x] = r; }
}
__global__ void __launch_bounds__(1024)
k_maxdiag_final(int Nc, const double* __restrict__ Hcc, const double* __restrict__ max_partial, int npartial, Scalars* __restrict__ sc) {
    __shared__ double smax[32];
    double m = 0.0, mp = 0.0;
    for (int j = threadIdx.x; j < 6 * Nc; j += blockDim.x) m = fmax(m, fabs(Hcc[36 * size_t(j / 6) + 7 * (j % 6)]));
    for (int j = threadIdx.x; j < npartial; j += blockDim.x) mp = fmax(mp, max_partial[j]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o)); mp = fmax(mp, __shfl_xor_sync(0xffffffffu, mp, o)); }
    if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) { double r = 0.0; for (int w = 0; w < 32; ++w) r = fmax(r, smax[w]); sc->maxdiag_f = r; }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = mp;
    __syncthreads();
    if (threadIdx.x == 0) { double r = 0.0; for (int w = 0; w < 32; ++w) r = fmax(r, smax[w]); sc->maxdiag_pt = r; }
}

// max |x - (x + (-g))| over free scalars — the minimizer's gradient_max_norm (Plus(x, -g) lifted back, SURVEY a-LM).
__global__ void __launch_bounds__(256)
k_gradmax_points(View V, const double* __restrict__ pt, double* __restrict__ max_partial) {
    __shared__ double smax[32];
    double m = 0.0;
    const int n = 3 * V.Np;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int p = i / 3;
        if (V.pt_off[p + 1] > V.pt_off[p]) { const double x = pt[i]; m = fmax(m, fabs(x - (x + (-V.gp[i])))); }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) { double r = 0.0; for (int w = 0; w < (blockDim.x >> 5); ++w) r = fmax(r, smax[w]); max_partial[blockIdx.x] = r; }
}

__global__ void __launch_bounds__(1024)
k_gradmax_final(int Nc, int Ng, int fi, const unsigned char* __restrict__ cam_free, const unsigned char* __restrict__ grp_free,
                const double* __restrict__ pose, const double* __restrict__ intr, const double* __restrict__ gc, const double* __restrict__ gi,
                const double* __restrict__ max_partial, int npartial, Scalars* __restrict__ sc) {
    __shared__ double smax[32];
    double m = 0.0, mp = 0.0;
    for (int j = threadIdx.x; j < 6 * Nc; j += blockDim.x) if (cam_free[j / 6]) { const double x = pose[j]; m = fmax(m, fabs(x - (x + (-gc[j])))); }
    if (fi) for (int j = threadIdx.x; j < 4 * Ng; j += blockDim.x) if (grp_free[j / 4]) { const double x = intr[j]; m = fmax(m, fabs(x - (x + (-gi[j])))); }
    for (int j = threadIdx.x; j < npartial; j += blockDim.x) mp = fmax(mp, max_partial[j]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o)); mp = fmax(mp, __shfl_xor_sync(0xffffffffu, mp, o)); }
    if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) { double r = 0.0; for (int w = 0; w < 32; ++w) r = fmax(r, smax[w]); sc->gmax_f = r; }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = mp;
    __syncthreads();
    if (threadIdx.x == 0) { double r = 0.0; for (int w = 0; w < 32; ++w) r = fmax(r, smax[w]); sc->gmax_pt = r; }
}

}  // namespace bak
