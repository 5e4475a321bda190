// ba_p2p.cuh — multi-GPU: the per-PCG-iteration all-reduce of the matvec result y (6 Nc doubles, 85 KB on Venice) as a ONE-SHOT
// peer-memory all-reduce over NVLink/NVSwitch, fused into the K5b gather and the PCG dot-product kernel ("B200-native step
// beyond NCCL").  NCCL's all-reduce of this payload is latency-bound (tens of microseconds, comparable to the local matvec at 8 GPUs).
//
//   k_camera_gather_push : K5b on every rank — sums camera c's partial rows (fixed order) into slot [parity] of the rank's OWN exchange
//                          buffer (local stores, no ticket, no flag).
//   k_pcg_tail (mode 2)  : (ba_pcg_cluster.cuh) first fences once at system scope and publishes flag[my rank] = seq on every rank
//                          (st.release.sys over NVLink; the kernel boundary already ordered the gather's stores), then
//                          waits (ld.acquire.sys on LOCAL memory) until all W flags show seq, then y = sum over ranks IN RANK ORDER of
//                          the W ranks' slots, read straight from the peers' memory (ld.relaxed.sys over NVLink; peer pointers mapped
//                          with CUDA IPC) — the same order on every rank, so all ranks hold bit-identical y and take identical
//                          decisions — and continues with q = y + Df2 p, p.q, alpha as k_pcg_q.
// (The first version PUSHED every camera's 6-vector into every peer's buffer; that needs a system-scope fence after the remote stores
// in each of the 1779 CTAs, which cost ~35 us per PCG iteration at 2 GPUs.  Pulling needs one.)
// Slots are double-buffered by the parity of seq: a rank can be at most one all-reduce ahead of the slowest rank (it needs that rank's
// slot of the current sequence number to finish), so parity p is never overwritten while someone still reads it.
// One rank per GPU only: the wait spins on flags another GPU writes (never emulate several ranks on one GPU with this path).
#pragma once
#include "ba_rows.cuh"

namespace bak {

constexpr int P2P_MAX_RANKS = 8;

struct PeerView {
    double* data[P2P_MAX_RANKS];        // rank r's exchange buffer: slot [parity] at data[r] + parity * ldf (data[rank] is the local one)
    unsigned int* flags[P2P_MAX_RANKS]; // rank r's flag array [W]: flags[r][s] = last sequence number rank s has fully delivered to r
    int W, rank, ldf;
    int* timeout_flag;                  // local Scalars::bad_lin: bit 8 set if a wait gave up (the host turns it into BA_ERR_NCCL)
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ double ld_relaxed_sys_f64(const double* p) {   // home-memory read (local HBM/L2 or a peer's, over NVLink); never a stale cached copy
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(GATHER_WARPS * 32, BA_GATHER_MINB)
k_camera_gather_push(int Nc, int Ng, int fi, int nTiles, RowSched S, const double* __restrict__ ylocal_rw, PeerView P, unsigned int seq,
                     PcgState2* __restrict__ st, int check_done) {
    __shared__ double sw[GATHER_WARPS][6];
    pdl_trigger();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int ioff = 6 * Nc;
    double* const mine = P.data[P.rank] + size_t(seq & 1u) * P.ldf;
    const int c = blockIdx.x;
    if (c >= Nc) { pdl_wait(); if (check_done && st->done) return; }
    if (c < Nc) {
        if (!camera_row_sum<true>(S, c, sw, check_done ? &st->done : nullptr)) return;
        if (threadIdx.x < 3) {   // thread h: 16 bytes of this camera's 6-vector
            const int h = threadIdx.x;
            const double2 v = make_double2(sw[0][2 * h], sw[0][2 * h + 1]);
            __stcg(reinterpret_cast<double2*>(mine + 6 * size_t(c) + 2 * h), v);
        }
    } else if (fi) {
        // intrinsics entries: single group -> ordered sum of the per-tile partials; several groups -> RED-accumulated in ylocal_rw already
        for (int k = wid; k < 4 * Ng; k += GATHER_WARPS) {
            double v;
            if (Ng == 1) {
                double s = 0.0;
                for (int t = lane; t < nTiles; t += 32) s += __ldcg(S.tile_intr + 4 * size_t(t) + k);
                v = warp_sum(s);
            } else v = ylocal_rw[ioff + k];
            if (lane == 0) __stcg(mine + ioff + k, v);
        }
    }
    // no ticket, no flag: the kernel boundary orders these stores before the tail kernel, which publishes the flags (ba_pcg_cluster.cuh)
}

// y = sum_r slot[r] in rank order (after all W ranks delivered sequence number seq); with do_dot: q = y + Df2 p in y, p.q, alpha.
__global__ void __launch_bounds__(128)
k_pcg_q_p2p(int nf, PeerView P, unsigned int seq, const double* __restrict__ Df2, const double* __restrict__ p, double* __restrict__ y,
            double* __restrict__ partial, PcgState2* __restrict__ st, int do_dot, int check_done) {
    __shared__ double sred[32];
    __shared__ bool last;
    if (check_done && st->done) return;
    if (threadIdx.x < P.W) {
        const unsigned int* f = P.flags[P.rank] + threadIdx.x;
        const long long t0 = clock64();
        while (int(ld_acquire_sys(f) - seq) < 0) {
            if ((*reinterpret_cast<volatile int*>(P.timeout_flag) & 8) || clock64() - t0 > 4000000000LL) { atomicOr(P.timeout_flag, 8); break; }   // ~2 s: a peer died; never hang the GPU
        }
    }
    __syncthreads();
    const size_t off = size_t(seq & 1u) * P.ldf;
    double acc = 0.0;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < nf; j += gridDim.x * blockDim.x) {
        double v[P2P_MAX_RANKS];
#pragma unroll
        for (int r = 0; r < P2P_MAX_RANKS; ++r) v[r] = r < P.W ? ld_relaxed_sys_f64(P.data[r] + off + j) : 0.0;   // all loads in flight, then
        double s = 0.0;
#pragma unroll
        for (int r = 0; r < P2P_MAX_RANKS; ++r) if (r < P.W) s += v[r];                                              // the sum in rank order
        if (do_dot) { const double pj = p[j]; s += Df2[j] * pj; acc += pj * s; }
        y[j] = s;
    }
    if (!do_dot) return;
    acc = block_sum(acc, sred);
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = acc;
        __threadfence();
        const unsigned int tk = atomicAdd(&st->ticket_r, 1u);
        last = (tk == gridDim.x - 1);
    }
    __syncthreads();
    double pq = 0.0;
    if (last) {
        __threadfence();
        for (int i = threadIdx.x; i < int(gridDim.x); i += blockDim.x) pq += __ldcg(partial + i);
        pq = block_sum(pq, sred);
    }
    if (last && threadIdx.x == 0) {
        if (!(pq > 0.0) || !isfinite(pq)) { st->done = 3; st->alpha = 0.0; }
        else st->alpha = st->rho / pq;
        st->ticket_r = 0;
    }
}

}  // namespace bak
