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
                     const PcgState2* __restrict__ st, int check_done) {
    __shared__ double sw[6];
    __shared__ int sidx[GATHER_IDX_CAP];
    pdl_trigger();
    double* const mine = P.data[P.rank] + size_t(seq & 1u) * P.ldf;
    const int c = blockIdx.x;
    if (c < Nc) {
        if (!camera_row_sum(S, c, sw, sidx, check_done ? &st->done : nullptr)) return;
        if (threadIdx.x < 3) {   // thread h: 16 bytes of this camera's 6-vector
            const int h = threadIdx.x;
            __stcg(reinterpret_cast<double2*>(mine + 6 * size_t(c) + 2 * h), make_double2(sw[2 * h], sw[2 * h + 1]));
        }
    } else {
        pdl_wait();
        if (check_done && st->done) return;
        if (fi) gather_intrinsics(Nc, Ng, nTiles, S, ylocal_rw, [&](int j, double v) { __stcg(mine + j, v); });
    }
    // no ticket, no flag: the kernel boundary orders these stores before the tail kernel, which publishes the flags (ba_pcg_cluster.cuh)
}

}  // namespace bak
