// bw_probe.cu — HBM ceilings relevant to the BA kernels: pure read, pure write, copy, and a 10-stream 16-byte-per-thread write
// (the store pattern of K1).  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bw_probe bw_probe.cu && ./bw_probe
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_read(const double2* __restrict__ a, size_t n, double* out) {
    double s = 0; for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) { double2 v = __ldg(a + i); s += v.x + v.y; }
    if (s == 1.2345) *out = s;
}
__global__ void k_write(double2* __restrict__ a, size_t n) {
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) a[i] = make_double2(1.0, 2.0);
}
__global__ void k_copy(const double2* __restrict__ a, double2* __restrict__ b, size_t n) {
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) b[i] = __ldg(a + i);
}
__global__ void k_write10(double2* __restrict__ a, size_t ld, size_t no) {   // one thread per observation, 10 column stores
    const size_t o = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (o < no) for (int k = 0; k < 10; ++k) a[k * ld + o] = make_double2(double(k), 2.0);
}
int main() {
    const size_t n = size_t(1) << 26;   // 1 GiB of double2
    double2 *a, *b; double* out;
    cudaMalloc(&a, n * 16); cudaMalloc(&b, n * 16); cudaMalloc(&out, 8);
    cudaMemset(a, 0, n * 16); cudaMemset(b, 0, n * 16);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto time = [&](const char* name, double bytes, auto f) {
        f(); f(); cudaDeviceSynchronize();
        cudaEventRecord(e0); for (int i = 0; i < 5; ++i) f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("%-28s %8.1f GB/s\n", name, bytes * 5 / (ms * 1e-3) / 1e9);
    };
    const int G = 148 * 8;
    time("read  (LDG.128)", n * 16.0, [&] { k_read<<<G, 512>>>(a, n, out); });
    time("write (STG.128)", n * 16.0, [&] { k_write<<<G, 512>>>(a, n); });
    time("copy  (r+w bytes)", n * 32.0, [&] { k_copy<<<G, 512>>>(a, b, n); });
    const size_t no = 5001946, ld = 5002016;
    time("write 10 columns x 16 B/thread", no * 160.0, [&] { k_write10<<<(no + 255) / 256, 256>>>(a, ld, no); });
    time("cudaMemset", n * 16.0, [&] { cudaMemsetAsync(a, 0, n * 16); });
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
