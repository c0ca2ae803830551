// ubench_poseidon.cu -- developer micro-benchmark (not part of the product): pure Poseidon permutation
// throughput of poseidon.cuh on one GPU, plus raw IMAD.WIDE / IADD3 issue rates for the pipe model in DESIGN.md.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_poseidon ubench_poseidon.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../../near-light-client_b200/csrc/poseidon.cuh"
typedef uint64_t u64; typedef uint32_t u32;

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

__global__ void __launch_bounds__(256) k_imad_wide(u64* io, int iters) {
    u64 a[8]; u32 x = threadIdx.x * 2654435761u + 1, y = blockIdx.x + 3;
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = io[threadIdx.x + i];
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = (u64)x * y + a[i];   // IMAD.WIDE.U32 with 64-bit accumulate
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = (u64)(u32)(a[i]>>7) * x + a[i];
    }
    u64 r = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) r ^= a[i];
    io[blockIdx.x * 256 + threadIdx.x] = r;
}
__global__ void __launch_bounds__(256) k_iadd3(u32* io, int iters) {
    u32 a[8]; u32 x = threadIdx.x * 2654435761u + 1, y = blockIdx.x + 3;
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = io[threadIdx.x + i];
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(x));
#pragma unroll
        for (int i = 0; i < 8; i++) asm volatile("xor.b32 %0, %0, %1;" : "+r"(a[i]) : "r"(y));
    }
    u32 r = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) r ^= a[i];
    io[blockIdx.x * 256 + threadIdx.x] = r;
}
__global__ void __launch_bounds__(256) k_mixed(u64* io, int iters) {
    u64 a[4]; u32 b[4]; u32 x = threadIdx.x * 2654435761u + 1, y = blockIdx.x + 3;
#pragma unroll
    for (int i = 0; i < 4; i++) { a[i] = io[threadIdx.x + i]; b[i] = (u32)io[threadIdx.x + 4 + i]; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
        for (int i = 0; i < 4; i++) { a[i] = (u64)(u32)(a[i]>>7) * x + a[i]; asm volatile("add.u32 %0, %0, %1;" : "+r"(b[i]) : "r"(y)); }
        }
    }
    u64 r = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) r ^= a[i] ^ b[i];
    io[blockIdx.x * 256 + threadIdx.x] = r;
}

template <class F> float time_ms(F f) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}

template <int THREADS, int MINB> void run_perm(u64* d, int sms) {
    const int iters = 32;
    for (int bps : {2, 4, 6, 8, 12, 16}) {
        if (bps * THREADS > 2048) continue;
        int blocks = sms * bps;
        float ms = time_ms([&] { k_perm<THREADS, MINB><<<blocks, THREADS>>>(d, iters); });
        cudaError_t e = cudaGetLastError();
        double perms = (double)blocks * THREADS * iters;
        printf("perm T=%d minb=%d blocks/SM=%d: %.3f ms  %.3f Gperm/s %s\n", THREADS, MINB, bps, ms, perms / ms * 1e-6, e ? cudaGetErrorString(e) : "");
    }
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    printf("%s SMs=%d clock=%d kHz\n", p.name, sms, p.clockRate);
    u64* d; cudaMalloc(&d, (size_t)sms * 2048 * 12 * 8 * 2); cudaMemset(d, 1, (size_t)sms * 2048 * 12 * 8 * 2);
    {
        int iters = 4096, blocks = sms * 8;
        float ms = time_ms([&] { k_imad_wide<<<blocks, 256>>>(d, iters); });
        double ops = (double)blocks * 256 * iters * 16;
        printf("IMAD.WIDE.U32 (64b acc): %.3f ms %.2f Tlane-op/s = %.1f lanes/clk/SM @%d MHz nominal\n", ms, ops / ms * 1e-9, ops / ms * 1e-3 / sms / (p.clockRate * 1e-3) , p.clockRate / 1000);
        ms = time_ms([&] { k_iadd3<<<blocks, 256>>>((u32*)d, iters); });
        printf("ADD/XOR u32: %.3f ms %.2f Tlane-op/s = %.1f lanes/clk/SM\n", ms, ops / ms * 1e-9, ops / ms * 1e-3 / sms / (p.clockRate * 1e-3));
        ms = time_ms([&] { k_mixed<<<blocks, 256>>>(d, iters); });
        double ops2 = (double)blocks * 256 * iters * 32;
        printf("mixed 1:1 IMAD.WIDE+ADD: %.3f ms %.2f Tlane-op/s = %.1f lanes/clk/SM\n", ms, ops2 / ms * 1e-9, ops2 / ms * 1e-3 / sms / (p.clockRate * 1e-3));
    }
    run_perm<128, 1>(d, sms);
    run_perm<128, 4>(d, sms);
    run_perm<128, 6>(d, sms);
    run_perm<256, 2>(d, sms);
    run_perm<64, 8>(d, sms);
    return 0;
}
