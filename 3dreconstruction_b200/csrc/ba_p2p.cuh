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
                     const PcgState2* __restrict__ st, int check_done, const double* __restrict__ G) {
    __shared__ double sw[6];
    __shared__ int sidx[GATHER_IDX_CAP];
    __shared__ double sG[36];
    pdl_trigger();
    double* const mine = P.data[P.rank] + size_t(seq & 1u) * P.ldf;
    const int c = blockIdx.x;
    if (c < Nc) {
        if (G && threadIdx.x < 36) sG[threadIdx.x] = __ldg(G + 36 * size_t(c) + threadIdx.x);
        if (!camera_row_sum(S, c, sw, sidx, check_done ? &st->done : nullptr)) return;
        if (threadIdx.x < 3) {   // thread h: 16 bytes of this camera's 6-vector (pose-factored rows: of G_c^T times the row sum; linear, so
            const int h = threadIdx.x;   // every rank applies it to its own partial before the all-reduce)
            const double lo = G ? pose_factor_apply_t(sG, sw, 2 * h) : sw[2 * h], hi = G ? pose_factor_apply_t(sG, sw, 2 * h + 1) : sw[2 * h + 1];
            __stcg(reinterpret_cast<double2*>(mine + 6 * size_t(c) + 2 * h), make_double2(lo, hi));
        }
    } else {
        pdl_wait();
        if (check_done && st->done) return;
        if (fi) gather_intrinsics(Nc, Ng, nTiles, S, ylocal_rw, [&](int j, double v) { __stcg(mine + j, v); });
    }
    // no ticket, no flag: the kernel boundary orders these stores before the tail kernel, which publishes the flags (ba_pcg_cluster.cuh)
}

// The pull of the one-shot all-reduce as a kernel of its own, on MANY SMs.  Inside the PCG tail (one cluster of 8 CTAs) the pull of W slots
// is bound by what 8 SMs can keep in flight over NVLink: 88 us per PCG iteration on the Final-13682 shape at 8 GPUs (W x 657 KB), the one
// term that does not shrink with the shard.  Here: block 0 fences once at system scope and publishes this rank's flag on every rank (the
// kernel boundary ordered the gather's stores), every block waits (local memory) until all W ranks have published this sequence number,
// then pulls ITS slice of all W slots (16-byte ld.relaxed.sys, all W loads of a thread in flight together), adds them in rank order (the same
// order on every rank: bit-identical y everywhere) and stores y locally.  The tail then runs in mode 1 (y complete in local memory).
// The grid is small enough to be co-resident (<= 2 blocks per SM), so the blocks that spin cannot keep block 0 from running.
constexpr int PULL_THREADS = 256;

__device__ __forceinline__ double2 ld_relaxed_sys_v2f64(const double* p) {
    double2 v;
    asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(PULL_THREADS)
k_allreduce_pull(int nf, PeerView P, unsigned int seq, double* __restrict__ y, const PcgState2* __restrict__ st, int check_done) {
    pdl_trigger();
    pdl_wait();
    if (check_done && st->done) return;
    if (blockIdx.x == 0 && threadIdx.x < P.W) { __threadfence_system(); st_release_sys(P.flags[threadIdx.x] + P.rank, seq); }
    if (threadIdx.x < P.W) {
        const unsigned int* f = P.flags[P.rank] + threadIdx.x;
        const long long t0 = clock64();
        while (int(ld_acquire_sys(f) - seq) < 0) {
            if ((*reinterpret_cast<volatile int*>(P.timeout_flag) & 8) || clock64() - t0 > 4000000000LL) { atomicOr(P.timeout_flag, 8); break; }   // ~2 s: a peer died; never hang the GPU
        }
    }
    __syncthreads();
    const size_t off = size_t(seq & 1u) * P.ldf;                 // ldf is a multiple of 16 doubles: slots and pairs are 16-byte aligned
    const int npair = (nf + 1) / 2;                              // the slot is padded to ldf >= nf + 1
    for (int i = blockIdx.x * PULL_THREADS + threadIdx.x; i < npair; i += gridDim.x * PULL_THREADS) {
        double2 v[P2P_MAX_RANKS];
#pragma unroll
        for (int r = 0; r < P2P_MAX_RANKS; ++r) v[r] = r < P.W ? ld_relaxed_sys_v2f64(P.data[r] + off + 2 * size_t(i)) : make_double2(0.0, 0.0);
        double2 s = make_double2(0.0, 0.0);
#pragma unroll
        for (int r = 0; r < P2P_MAX_RANKS; ++r) if (r < P.W) { s.x += v[r].x; s.y += v[r].y; }
        if (2 * i + 1 < nf) *reinterpret_cast<double2*>(y + 2 * size_t(i)) = s;
        else y[2 * size_t(i)] = s.x;
    }
}

}  // namespace bak
