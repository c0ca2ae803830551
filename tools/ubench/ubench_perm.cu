// ubench_perm.cu -- developer micro-benchmark: pure Poseidon permutation throughput of csrc/poseidon.cuh.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a [-DNLC_ZMASK=k] -o ubench_perm ubench_perm.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../near-light-client_b200/csrc/poseidon.cuh"
typedef uint64_t u64;
template <int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_perm(u64* io, int iters) {
    u64 s[12];
    const size_t t = blockIdx.x * (size_t)THREADS + threadIdx.x;
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = io[t * 12 + i];
    for (int it = 0; it < iters; it++) poseidon::permute_nc(s);
#pragma unroll
    for (int i = 0; i < 12; i++) io[t * 12 + i] = gl::canon(s[i]);
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    u64* d; cudaMalloc(&d, (size_t)sms * 2048 * 12 * 8); cudaMemset(d, 1, (size_t)sms * 2048 * 12 * 8);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    auto run = [&](auto kern, int minb, int bps) {
        int blocks = sms * bps, iters = 64;
        kern<<<blocks, 128>>>(d, iters); cudaDeviceSynchronize();
        cudaEventRecord(a); kern<<<blocks, 128>>>(d, iters); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        printf("ZMASK=%d minb=%d blocks/SM=%d: %.3f ms %.3f Gperm/s\n", NLC_ZMASK, minb, bps, ms, (double)blocks * 128 * iters / ms * 1e-6);
    };
    run(k_perm<128, 4>, 4, 4); run(k_perm<128, 5>, 5, 5); run(k_perm<128, 6>, 6, 6); run(k_perm<128, 8>, 8, 8);
    return 0;
}
