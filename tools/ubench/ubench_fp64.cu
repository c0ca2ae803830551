// ubench_fp64.cu -- is the FP64 pipe of this GPU a usable third integer-capable pipe?  DFMA / DADD issue rate per SM
// (exact integer arithmetic below 2^53), alone and interleaved with IMAD / IADD3 streams.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_fp64 ubench_fp64.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int MODE>
__global__ void __launch_bounds__(256) k(double* out, uint32_t* iout, int iters, double seed, uint32_t iseed) {
    double a0 = seed, a1 = seed + 1, a2 = seed + 2, a3 = seed + 3, a4 = seed + 4, a5 = seed + 5, a6 = seed + 6, a7 = seed + 7;
    uint32_t i0 = iseed, i1 = iseed + 1, i2 = iseed + 2, i3 = iseed + 3, i4 = iseed + 4, i5 = iseed + 5, i6 = iseed + 6, i7 = iseed + 7;
    const double c = seed * 0.5, d = 3.0;
    for (int it = 0; it < iters; it++) {
        if (MODE == 0 || MODE == 2 || MODE == 3) { // 8 independent DFMA
            a0 = fma(a0, c, d); a1 = fma(a1, c, d); a2 = fma(a2, c, d); a3 = fma(a3, c, d);
            a4 = fma(a4, c, d); a5 = fma(a5, c, d); a6 = fma(a6, c, d); a7 = fma(a7, c, d);
        }
        if (MODE == 1) { // 8 independent DADD
            a0 += d; a1 += d; a2 += d; a3 += d; a4 += d; a5 += d; a6 += d; a7 += d;
        }
        if (MODE == 2 || MODE == 4) { // 8 independent 32-bit IMAD (FMA-heavy pipe)
            i0 = i0 * iseed + 7; i1 = i1 * iseed + 7; i2 = i2 * iseed + 7; i3 = i3 * iseed + 7;
            i4 = i4 * iseed + 7; i5 = i5 * iseed + 7; i6 = i6 * iseed + 7; i7 = i7 * iseed + 7;
        }
        if (MODE == 3 || MODE == 5) { // 8 independent LOP3 / shifts (ALU pipe)
            i0 = (i0 ^ iseed) >> 1 | i0 << 3; i1 = (i1 ^ iseed) >> 1 | i1 << 3; i2 = (i2 ^ iseed) >> 1 | i2 << 3; i3 = (i3 ^ iseed) >> 1 | i3 << 3;
            i4 = (i4 ^ iseed) >> 1 | i4 << 3; i5 = (i5 ^ iseed) >> 1 | i5 << 3; i6 = (i6 ^ iseed) >> 1 | i6 << 3; i7 = (i7 ^ iseed) >> 1 | i7 << 3;
        }
    }
    out[blockIdx.x * 256 + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    iout[blockIdx.x * 256 + threadIdx.x] = i0 ^ i1 ^ i2 ^ i3 ^ i4 ^ i5 ^ i6 ^ i7;
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, blocks = sms * 8, iters = 1 << 14;
    double* d; uint32_t* di; cudaMalloc(&d, blocks * 256 * 8); cudaMalloc(&di, blocks * 256 * 4);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    auto run = [&](auto kern, const char* name, double ops_per_iter) {
        kern<<<blocks, 256>>>(d, di, iters, 1.0000001, 12345u); cudaDeviceSynchronize();
        cudaEventRecord(a); kern<<<blocks, 256>>>(d, di, iters, 1.0000001, 12345u); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        const double clk = p.clockRate * 1e3; // Hz
        const double thread_ops = (double)blocks * 256 * iters * ops_per_iter;
        printf("%-34s %8.3f ms  %7.1f thread-ops/clk/SM  (%.2f T ops/s)\n", name, ms, thread_ops / (ms * 1e-3) / clk / sms, thread_ops / (ms * 1e-3) * 1e-12);
    };
    run(k<0>, "DFMA x8", 8); run(k<1>, "DADD x8", 8); run(k<4>, "IMAD x8", 8); run(k<5>, "ALU (LOP3/SHF) x8 stmts", 8);
    run(k<2>, "DFMA x8 + IMAD x8", 16); run(k<3>, "DFMA x8 + ALU x8 stmts", 16);
    return 0;
}
