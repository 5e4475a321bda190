// chol_probe.cu — how long does one warp take to factorise + invert a 32x32 tile (ba_dense.cuh::potrf_trinv32), and what do a tile load +
// 32x32x32 product cost?  Single CTA, clock64 around each phase.   nvcc -arch=sm_100a -O3 -I../../3dreconstruction_b200/csrc chol_probe.cu
#include <cstdio>
#include <vector>
#include "ba_dense.cuh"
using namespace bak;

__global__ void __launch_bounds__(256) probe(const double* A, double* out, long long* cyc, int reps) {
    __shared__ double sL[DT * DTP], sZt[DT * DTP], sA[DT * DTP];
    __shared__ __align__(16) double sLb[DT * DTB];
    __shared__ __align__(16) double scol[DT];
    __shared__ double srinv[DT];
    const int tid = threadIdx.x, lane = tid & 31;
    long long t_fac = 0, t_mma = 0, t_load = 0;
    double acc[4] = {0, 0, 0, 0};
    for (int it = 0; it < reps; ++it) {
        __syncthreads();
        long long t0 = clock64();
        tile_load_cg(A, DT, sL, tid);
        tile_load_cg(A, DT, sA, tid);
        __syncthreads();
        long long t1 = clock64();
        if (tid < 32) potrf_trinv32(sL, sZt, sLb, scol, srinv, lane);
        __syncthreads();
        long long t2 = clock64();
        double o[4];
        tile_mma(sA, sZt, tid, o);
        acc[0] += o[0]; acc[1] += o[1]; acc[2] += o[2]; acc[3] += o[3];
        __syncthreads();
        long long t3 = clock64();
        t_load += t1 - t0; t_fac += t2 - t1; t_mma += t3 - t2;
    }
    if (tid == 0) { cyc[0] = t_load / reps; cyc[1] = t_fac / reps; cyc[2] = t_mma / reps; }
    out[tid] = acc[0] + acc[1] + acc[2] + acc[3] + sZt[tid];
}

int main() {
    std::vector<double> h(DT * DT, 0.0);
    for (int i = 0; i < DT; ++i) for (int j = 0; j <= i; ++j) h[i * DT + j] = (i == j) ? 40.0 + i : 1.0 / (1 + i + j);
    double *A, *out; long long* cyc;
    cudaMalloc(&A, sizeof(double) * DT * DT); cudaMalloc(&out, sizeof(double) * 256); cudaMalloc(&cyc, 3 * sizeof(long long));
    cudaMemcpy(A, h.data(), sizeof(double) * DT * DT, cudaMemcpyHostToDevice);
    probe<<<1, 256>>>(A, out, cyc, 200);
    long long c[3];
    cudaMemcpy(c, cyc, sizeof c, cudaMemcpyDeviceToHost);
    int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    printf("cycles: load %lld  potrf+trinv %lld  mma %lld   (clock %d MHz: %.2f / %.2f / %.2f us)  err=%s\n", c[0], c[1], c[2], khz / 1000, c[0] * 1e3 / khz, c[1] * 1e3 / khz, c[2] * 1e3 / khz,
           cudaGetErrorString(cudaGetLastError()));
    // check: Zt * L^T should be identity -> verify on host the inverse of the first column
    return 0;
}
