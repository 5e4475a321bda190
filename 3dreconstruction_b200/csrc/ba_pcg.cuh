// ba_pcg.cuh — the start of a PCG solve (k_pcg_init2); everything an iteration does after the Schur product is in the cluster tail
// (ba_pcg_cluster.cuh).
#pragma once
#include "ba_kernels.cuh"

namespace bak {

constexpr int PCG_BLOCK = 128;
#ifndef BA_K5F_VT256
#define BA_K5F_VT256 0         // 1: vt records padded to 64 bytes and fetched with two 32-byte loads (measured 3 % slower than three 16-byte loads of packed records)
#endif
constexpr int VT_STRIDE = BA_K5F_VT256 ? 8 : 6;   // doubles per camera in vt = G p

// PCG start: b = reduced rhs (already reduced into `b`), x = 0, r = b, z = Minv b, p = z, rho = b.z (single CTA, once per LM step).
// (p is __restrict__ but read back after the barrier only through the same pointer.)
__global__ void __launch_bounds__(1024)
k_pcg_init2(int nf, int Nc, const double* __restrict__ Minv, const double* __restrict__ b, double* __restrict__ x,
            double* __restrict__ r, double* __restrict__ z, double* __restrict__ p, PcgState2* __restrict__ st,
            const double* __restrict__ G, double* __restrict__ vt) {
    __shared__ double sred[32];
    double acc = 0.0;
    for (int j = threadIdx.x; j < nf; j += blockDim.x) {
        x[j] = 0.0;
        const double zj = apply_minv(j, Nc, Minv, b);
        r[j] = b[j]; z[j] = zj; p[j] = zj;
        acc += b[j] * zj;
    }
    if (vt) {   // pose-factored matvec: vt = [G_c p_c | p of the intrinsics entries] (ba_rows.cuh: k_pose_factors)
        __syncthreads();
        for (int j = threadIdx.x; j < nf; j += blockDim.x) {
            if (j < 6 * Nc) {
                const int c = j / 6, a = j - 6 * c;
                const double* Gr = G + 36 * size_t(c) + 6 * a;
                const double* pc = p + 6 * c;
                double v = 0.0;
#pragma unroll
                for (int k = 0; k < 6; ++k) v += Gr[k] * pc[k];
                vt[VT_STRIDE * size_t(c) + a] = v;
            } else vt[j + (VT_STRIDE - 6) * Nc] = p[j];
        }
    }
    acc = block_sum(acc, sred);
    if (threadIdx.x == 0) {
        st->rho = acc; st->rho0 = acc; st->rel = acc > 0.0 ? 1.0 : 0.0; st->Q0 = 0.0; st->alpha = 0.0; st->beta = 0.0; st->k = 0;
        st->done = (acc > 0.0 && isfinite(acc)) ? 0 : (acc == 0.0 ? 1 : 3);
        st->ticket_q = 0; st->ticket_r = 0;
    }
}

}  // namespace bak
