// ba_pcg.cuh — the start of a PCG solve (k_pcg_init2); everything an iteration does after the Schur product is in the cluster tail
// (ba_pcg_cluster.cuh).
#pragma once
#include "ba_kernels.cuh"

namespace bak {

constexpr int PCG_BLOCK = 128;

// PCG start: b = reduced rhs (already reduced into `b`), x = 0, r = b, z = Minv b, p = z, rho = b.z (single CTA, once per LM step).
__global__ void __launch_bounds__(1024)
k_pcg_init2(int nf, int Nc, const double* __restrict__ Minv, const double* __restrict__ b, double* __restrict__ x,
            double* __restrict__ r, double* __restrict__ z, double* __restrict__ p, PcgState2* __restrict__ st) {
    __shared__ double sred[32];
    double acc = 0.0;
    for (int j = threadIdx.x; j < nf; j += blockDim.x) {
        x[j] = 0.0;
        const double zj = apply_minv(j, Nc, Minv, b);
        r[j] = b[j]; z[j] = zj; p[j] = zj;
        acc += b[j] * zj;
    }
    acc = block_sum(acc, sred);
    if (threadIdx.x == 0) {
        st->rho = acc; st->rho0 = acc; st->rel = acc > 0.0 ? 1.0 : 0.0; st->Q0 = 0.0; st->alpha = 0.0; st->beta = 0.0; st->k = 0;
        st->done = (acc > 0.0 && isfinite(acc)) ? 0 : (acc == 0.0 ? 1 : 3);
        st->ticket_q = 0; st->ticket_r = 0;
    }
}

}  // namespace bak
