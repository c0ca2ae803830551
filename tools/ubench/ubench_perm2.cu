// ubench_perm2.cu -- developer micro-benchmark: Poseidon permutation throughput of csrc/poseidon.cuh with the warps
// of an SM IN PHASE (all start together, as a static deal leaves them) against OUT OF PHASE (every warp spins a
// pseudo-random part of one permutation's duration first, as claimed work units leave them after a few units).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_perm2 ubench_perm2.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../near-light-client_b200/csrc/poseidon.cuh"
typedef uint64_t u64;
template <int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_perm(u64* io, int iters, int stagger) {
    u64 s[12];
    const size_t t = blockIdx.x * (size_t)THREADS + threadIdx.x;
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = io[t * 12 + i];
    if (stagger) {
        const unsigned w = (unsigned)(t >> 5) * 2654435761u;
        const long long until = clock64() + (long long)((w >> 8) % (unsigned)stagger);
        while (clock64() < until) {}
    }
    for (int it = 0; it < iters; it++) poseidon::permute_nc(s);
#pragma unroll
    for (int i = 0; i < 12; i++) io[t * 12 + i] = gl::canon(s[i]);
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    u64* d; cudaMalloc(&d, (size_t)sms * 4096 * 12 * 8); cudaMemset(d, 1, (size_t)sms * 4096 * 12 * 8);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    auto run = [&](auto kern, int threads, int minb, int bps, int stagger) {
        int blocks = sms * bps, iters = 128;
        kern<<<blocks, threads>>>(d, iters, stagger); cudaDeviceSynchronize();
        cudaEventRecord(a); kern<<<blocks, threads>>>(d, iters, stagger); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        printf("threads=%d minb=%d blocks/SM=%d stagger=%6d clk: %.3f ms %.3f Gperm/s\n", threads, minb, bps, stagger, ms,
               (double)blocks * threads * iters / ms * 1e-6);
    };
    for (int st : {0, 40000}) {
        run(k_perm<128, 3>, 128, 3, 3, st); run(k_perm<128, 4>, 128, 4, 4, st); run(k_perm<128, 5>, 128, 5, 5, st);
        run(k_perm<128, 6>, 128, 6, 6, st); run(k_perm<128, 8>, 128, 8, 8, st);
        run(k_perm<64, 10>, 64, 10, 10, st); run(k_perm<32, 20>, 32, 20, 20, st); run(k_perm<32, 24>, 32, 24, 24, st);
    }
    return 0;
}
