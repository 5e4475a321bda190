// ba_stream.cuh — K4/K5a as a persistent, warp-specialised streaming kernel (scatter_mode 6, an experimental opt-in; see ba_engine.cu: do_set_problem for why it is not the default).
//
// Same tile math, same static schedule and bit-identical partial rows as k_schur_rows (ba_rows.cuh); what changes is how the bytes
// arrive.  ncu on the CTA-per-tile shape shows it latency-exposed, not bandwidth-bound (0.68 of the measured HBM peak at 35 % warps
// active, long_scoreboard 10 per issue): a CTA has bytes in flight only during the first quarter of its life, so an SM sustains
// ~55 KB of outstanding loads where ~45 KB is the bare minimum for the HBM rate.  Here:
//   * one CTA per SM, 1 producer warp + NG consumer groups of 128 threads (one tile each at a time);
//   * the producer streams tiles t = blockIdx.x, + gridDim.x, ... through a ring of NST shared-memory stages with cp.async.bulk
//     (TMA 1-D: one copy per stored Jacobian column piece of the tile, one for its schedule records, one for the Hinv blocks of its
//     points or its residuals), completion on an mbarrier per stage (expect_tx = the tile's bytes);
//   * a consumer group waits for its stage, copies its operands into registers and hands the stage back AT ONCE (mbarrier arrive), so a
//     stage is occupied for ~100 ns and practically the whole ring (8 x 21 KB on the Venice configuration) is bytes in flight,
//     independent of what the consumers are doing; the tile math then runs from registers with group-local named barriers;
//   * everything the producer loads is static during a linear solve, so under programmatic dependent launch it fills the ring before
//     griddepcontrol.wait; the consumers wait, read the stop flag, and either start or leave (then the producer drains its copies).
#pragma once
#include "ba_rows.cuh"

namespace bak {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (TMA 1-D); bytes and both addresses are multiples of 16; completion is signalled on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
                 "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ int ld_volatile_s32(const int* p) {
    int v;
    asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

#ifndef BA_STREAM_GROUPS
#define BA_STREAM_GROUPS 4          // consumer groups per CTA (A/B knob: 4 = 17 warps, 6 = 25 warps at <= 80 registers)
#endif
#ifndef BA_STREAM_SPARE_KB
#define BA_STREAM_SPARE_KB 0        // shared memory left free per SM (A/B knob: room for the next kernel's CTAs to run their prologue early)
#endif

template <int MODE, bool FI, int PK>
struct StreamCfg {
    static constexpr int NG = BA_STREAM_GROUPS;
    static constexpr int NCOL = (FI ? 13 : 9) - PK;                      // stored Jacobian column pairs
    static constexpr int J_BYTES = NCOL * TILE_OBS * 16;
    static constexpr int META_OFF = J_BYTES;                              // int4[TILE_OBS]
    static constexpr int AUX_OFF = META_OFF + TILE_OBS * 16;              // MATVEC: Hinv[TILE_MAX_PTS][6]; RHS: r2[TILE_OBS]
    static constexpr int AUX_BYTES = MODE == SWEEP_MATVEC ? TILE_MAX_PTS * 48 : TILE_OBS * 16;
    static constexpr int STAGE_BYTES = AUX_OFF + AUX_BYTES;               // a multiple of 128
    static constexpr int THREADS = (NG * 4 + 1) * 32;
    static constexpr int SCR_STRIDE = (int(sizeof(TileScratch)) + 127) / 128 * 128;
    static constexpr int BAR_BYTES = 256;
    static constexpr int BUDGET = (227 - BA_STREAM_SPARE_KB) * 1024 - BAR_BYTES - NG * SCR_STRIDE;
    static constexpr int NST = BUDGET / STAGE_BYTES < 12 ? BUDGET / STAGE_BYTES : 12;
    static constexpr int SMEM = BAR_BYTES + NG * SCR_STRIDE + NST * STAGE_BYTES;
    static_assert(NST >= 2, "the ring needs at least two stages");
    static_assert(STAGE_BYTES % 128 == 0, "stage alignment");
};

template <int MODE, bool FI, int PK>
__global__ void __launch_bounds__(StreamCfg<MODE, FI, PK>::THREADS, 1)
k_schur_stream(View V, RowSched S, const double* __restrict__ xin, double* __restrict__ atomic_out, const int* __restrict__ done_flag) {
    using Cfg = StreamCfg<MODE, FI, PK>;
    constexpr int NG = Cfg::NG, NST = Cfg::NST, NCOL = Cfg::NCOL;
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);            // [NST] producer -> consumers (transaction count)
    uint64_t* empty = full + NST;                                  // [NST] consumers -> producer (one arrival per warp of the consuming group)
    unsigned char* scratch0 = smem + Cfg::BAR_BYTES;
    unsigned char* stage0 = scratch0 + NG * Cfg::SCR_STRIDE;

    pdl_trigger();
    if (threadIdx.x == 0) {
        for (int s = 0; s < NST; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, TILE_OBS / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int G = gridDim.x, nT = V.nTiles;
    const int my = int(blockIdx.x) < nT ? (nT - 1 - int(blockIdx.x)) / G + 1 : 0;   // tiles of this CTA: blockIdx.x + i G, i < my
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == NG * 4) {
        // ------------------------------------------------ producer warp ------------------------------------------------
        bool waited = false;
        const size_t ld = V.ldo;
        for (int i0 = 0; i0 < my; i0 += 32) {
            int4 til = make_int4(0, 0, 0, 0);
            if (i0 + lane < my) til = __ldg(S.tile_info + blockIdx.x + size_t(i0 + lane) * G);
            const int lim = min(32, my - i0);
            for (int j = 0; j < lim; ++j) {
                const int i = i0 + j, s = i % NST, use = i / NST;
                const int4 ti = make_int4(__shfl_sync(0xffffffffu, til.x, j), __shfl_sync(0xffffffffu, til.y, j), __shfl_sync(0xffffffffu, til.z, j),
                                          __shfl_sync(0xffffffffu, til.w, j));
                if (i == NST && !waited) {     // the first ring fill needs nothing from the predecessor grid; refills need the consumers, who do
                    waited = true;
                    pdl_wait();
                    if (done_flag && ld_volatile_s32(done_flag)) {   // PCG already stopped: the consumers have left; let the copies land, then leave
                        if (lane == 0) for (int q = 0; q < NST; ++q) mbar_wait(full + q, 0);
                        return;
                    }
                }
                if (lane == 0) {
                    if (use > 0) mbar_wait(empty + s, unsigned(use - 1) & 1u);
                    const int o0 = ti.x, n = ti.y & 0xffffff, npts = int(unsigned(ti.y) >> 24), pfirst = ti.w;
                    if (n > TILE_OBS) mbar_arrive(full + s);          // long-point tile: the consumers read it from global memory
                    else {
                        unsigned char* st = stage0 + size_t(s) * Cfg::STAGE_BYTES;
                        const unsigned colb = unsigned(n) * 16u;
                        const unsigned aux = MODE == SWEEP_MATVEC ? unsigned(npts) * 48u : colb;
                        mbar_arrive_expect_tx(full + s, unsigned(NCOL + 1) * colb + aux);
                        const double2* J = V.J2 + o0;
#pragma unroll
                        for (int k = 0; k < NCOL; ++k) bulk_g2s(st + k * (TILE_OBS * 16), J + k * ld, colb, full + s);
                        bulk_g2s(st + Cfg::META_OFF, S.tmeta + o0, colb, full + s);
                        if (MODE == SWEEP_MATVEC) bulk_g2s(st + Cfg::AUX_OFF, V.Hinv + 6 * size_t(pfirst), aux, full + s);
                        else bulk_g2s(st + Cfg::AUX_OFF, V.r2 + o0, aux, full + s);
                    }
                }
            }
        }
        if (!waited) {   // fewer tiles than stages: still honour the stop flag before leaving shared memory behind
            pdl_wait();
            if (done_flag && ld_volatile_s32(done_flag)) { if (lane == 0) for (int q = 0; q < min(my, NST); ++q) mbar_wait(full + q, 0); }
        }
        return;
    }

    // ------------------------------------------------ consumer groups ------------------------------------------------
    const int grp = warp >> 2, tid = threadIdx.x & (TILE_OBS - 1);
    TileScratch& M = *reinterpret_cast<TileScratch*>(scratch0 + size_t(grp) * Cfg::SCR_STRIDE);
    const BarGroup bar{1 + grp};
    pdl_wait();
    if (done_flag && ld_volatile_s32(done_flag)) return;
    // one arrival per warp hands a stage back: the warp's shared-memory reads and its lane 0's mbarrier arrive go down the same in-order
    // pipe (128 same-address arrivals per tile cost more than the copies they protect)
    auto release = [&](int s) { __syncwarp(); if (lane == 0) mbar_arrive(empty + s); };
    int4 ti_next = grp < my ? __ldg(S.tile_info + blockIdx.x + size_t(grp) * G) : make_int4(0, 0, 0, 0);
    for (int i = grp; i < my; i += NG) {
        const int t = blockIdx.x + i * G;
        const int s = i % NST, use = i / NST;
        const int4 ti = ti_next;
        if (i + NG < my) ti_next = __ldg(S.tile_info + t + size_t(NG) * G);      // the next tile's record is in flight during this tile
        const int o0 = ti.x, n = ti.y & 0xffffff, hbase = ti.z, pfirst = ti.w;
        double yi[4] = {0, 0, 0, 0};
        mbar_wait(full + s, unsigned(use) & 1u);
        if (n <= TILE_OBS) {
            const unsigned char* st = stage0 + size_t(s) * Cfg::STAGE_BYTES;
            TileRegs<FI> R;
            R.tm = make_int4(0, 0, -1, 0);
            R.rr = make_double2(0.0, 0.0);
            R.h0 = R.h1 = R.h2 = make_double2(0.0, 0.0);
            if (tid < n) {
                const double2* col = reinterpret_cast<const double2*>(st) + tid;
                R.tm = reinterpret_cast<const int4*>(st + Cfg::META_OFF)[tid];
                R.jc[0] = col[0]; R.jc[1] = col[TILE_OBS]; R.jc[2] = col[2 * TILE_OBS];
                const double2 d = col[3 * TILE_OBS];
                if (PK) { R.jc[3] = make_double2(d.x, 0.0); R.jc[4] = make_double2(0.0, d.y); R.jc[5] = col[4 * TILE_OBS]; }
                else { R.jc[3] = d; R.jc[4] = col[4 * TILE_OBS]; R.jc[5] = col[5 * TILE_OBS]; }
#pragma unroll
                for (int k = 0; k < 3; ++k) R.jp[k] = col[(6 - PK + k) * TILE_OBS];
                if (FI) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) R.ji[k] = col[(9 - PK + k) * TILE_OBS];
                }
                if (MODE == SWEEP_RHS) R.rr = reinterpret_cast<const double2*>(st + Cfg::AUX_OFF)[tid];
            }
            if (MODE == SWEEP_MATVEC && (R.tm.w >> 16) > 0) {
                const double2* H = reinterpret_cast<const double2*>(st + Cfg::AUX_OFF) + 3 * tid;
                R.h0 = H[0]; R.h1 = H[1]; R.h2 = H[2];
            }
            // The stage may only be handed back once the shared-memory reads above have RETURNED: ptxas puts no scoreboard wait between the
            // LDS instructions and the mbarrier arrive (nothing in the arrive reads their destination registers), so an arrive issued
            // right behind them can overtake them in the memory pipe and the producer's next bulk copy lands on bytes that are still being
            // read (seen as PCG runs that differ from launch to launch on the Final-13682 shape).  A branch on a value formed from the last
            // loads of every kind makes the warp wait for them first; shared-memory loads of one warp return in order.
            {
                const double touch = R.jp[2].y + R.jc[5].y + R.h2.y + R.rr.y + double(R.tm.w);
                if (touch == -1.2345678901234567e+301) atomicAdd(&S.rows[0], touch);      // never taken; keeps the dependency alive
            }
            release(s);                                // operands are in registers: the stage can be refilled
            tile_rows<MODE, FI>(V, S, M, bar, tid, n, hbase, pfirst, R, xin, atomic_out, yi);
        } else {
            release(s);
            tile_rows_long<MODE, FI, PK>(V, S, M.sred, bar, tid, o0, n, hbase, pfirst, xin, atomic_out, yi);
        }
        if (FI && V.Ng == 1) tile_intr_store(S, M.sred, bar, tid, t, yi);
        // The next tile of this group may overwrite sC (RHS: right away) or sW/sred while slower threads still read them
        if (MODE == SWEEP_RHS || n > TILE_OBS) bar();
    }
}

}  // namespace bak
