// ubench_warp.cu -- developer micro-benchmark: latency of ONE transcript::warp_permute (single warp on the GPU).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_warp ubench_warp.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../near-light-client_b200/csrc/transcript.cuh"
__global__ void k_lat(u64* io, int iters, long long* cyc) {
    const int lane = threadIdx.x & 31;
    u64 x = io[lane];
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) x = transcript::warp_permute(x, lane);
    long long t1 = clock64();
    io[lane] = x;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_sbox(u64* io, int iters, long long* cyc) {
    u64 x = io[threadIdx.x];
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) x = poseidon::sbox7(x);
    long long t1 = clock64();
    io[threadIdx.x] = x;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    u64* d; long long* c; cudaMalloc(&d, 32 * 8); cudaMalloc(&c, 8); cudaMemset(d, 1, 32 * 8);
    long long h;
    for (int rep = 0; rep < 2; rep++) {
        k_lat<<<1, 32>>>(d, 100, c); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("warp_permute: %.0f cycles per permutation (%.1f per round)\n", h / 100.0, h / 3000.0);
        k_sbox<<<1, 32>>>(d, 1000, c); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("sbox7 dependent chain: %.1f cycles\n", h / 1000.0);
    }
    return 0;
}
