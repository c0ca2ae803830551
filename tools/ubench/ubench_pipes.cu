// ubench_pipes.cu -- developer micro-benchmark (not part of the product): issue rates of the integer / fp32
// instructions the Goldilocks kernels are built from, alone and paired, in lanes/clk/SM.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_pipes ubench_pipes.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64; typedef unsigned int u32;

#define REP8(X) X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7)

template <int KIND>
__global__ void __launch_bounds__(256) k(u32* io, int iters) {
    u32 a[8], b[8]; u64 w[8]; float f[8], g[8];
    u32 x = io[threadIdx.x] | 1, y = io[threadIdx.x + 1] | 3;
    float fx = __int_as_float((x & 0xffff) | 0x3f800000), fy = 0.25f;
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = io[threadIdx.x + i + 2]; b[i] = a[i] ^ x; w[i] = ((u64)a[i] << 32) | b[i]; f[i] = (float)(a[i] & 0xff); g[i] = f[i] + 1.0f; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int rr = 0; rr < 4; rr++) {
#define IMADLO(i)  asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(x), "r"(y));
#define IMADHI(i)  asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(x), "r"(y));
#define IMADW(i)   asm volatile("{.reg .u32 t; cvt.u32.u64 t, %0; mad.wide.u32 %0, t, %1, %0;}" : "+l"(w[i]) : "r"(x));
#define IMADWD(i)  asm volatile("{.reg .u32 t; cvt.u32.u64 t, %0; mad.wide.u32 %0, t, %1, %0;}" : "+l"(w[i]) : "r"(x));
#define MULW(i)    asm volatile("{.reg .u32 t; cvt.u32.u64 t, %0; mul.wide.u32 %0, t, %1;}" : "+l"(w[i]) : "r"(x));
#define IADD(i)    asm volatile("add.u32 %0, %0, %1;" : "+r"(b[i]) : "r"(y));
#define IADD3_(i)  asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(b[i]) : "r"(y), "r"(x));
#define LOP(i)     asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(y), "r"(x));
#define SHFL_(i)   asm volatile("shf.l.wrap.b32 %0, %0, %1, 5;" : "+r"(b[i]) : "r"(y));
#define LEA_(i)    asm volatile("{.reg .u32 t; shl.b32 t, %0, 3; add.u32 %0, t, %1;}" : "+r"(b[i]) : "r"(y));
#define PRMT_(i)   asm volatile("prmt.b32 %0, %0, %1, 0x5140;" : "+r"(b[i]) : "r"(y));
#define FFMA_(i)   asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(fx), "f"(fy));
#define FFMA2_(i)  asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(g[i]) : "f"(fx), "f"(fy));
#define ADDCC(i)   asm volatile("add.cc.u32 %0, %0, %2; addc.u32 %1, %1, %3;" : "+r"(a[i]), "+r"(b[i]) : "r"(x), "r"(y));
            if (KIND == 0) { REP8(IMADLO) }
            if (KIND == 1) { REP8(IMADHI) }
            if (KIND == 2) { REP8(IMADW) }
            if (KIND == 3) { REP8(IADD) }
            if (KIND == 4) { REP8(LOP) }
            if (KIND == 5) { REP8(SHFL_) }
            if (KIND == 6) { REP8(LEA_) }
            if (KIND == 7) { REP8(PRMT_) }
            if (KIND == 8) { REP8(FFMA_) }
            if (KIND == 9) { REP8(IMADW) REP8(IADD) }
            if (KIND == 10) { REP8(IMADW) REP8(IADD) REP8(LOP) }
            if (KIND == 11) { REP8(IMADLO) REP8(IADD) }
            if (KIND == 12) { REP8(FFMA_) REP8(IADD) }
            if (KIND == 13) { REP8(FFMA_) REP8(IMADLO) }
            if (KIND == 14) { REP8(FFMA_) REP8(IMADW) }
            if (KIND == 15) { REP8(ADDCC) }
            if (KIND == 16) { REP8(IADD3_) }
            if (KIND == 17) { REP8(FFMA_) REP8(FFMA2_) REP8(IADD) REP8(LOP) }
            if (KIND == 18) { REP8(IMADW) REP8(IADD) REP8(LOP) REP8(SHFL_) REP8(PRMT_) }
            if (KIND == 19) { REP8(MULW) REP8(IADD) }
        }
    }
    u32 r = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) r ^= a[i] ^ b[i] ^ (u32)w[i] ^ (u32)(w[i] >> 32) ^ __float_as_uint(f[i]) ^ __float_as_uint(g[i]);
    io[blockIdx.x * 256 + threadIdx.x] = r;
}

static const char* NAMES[] = {"IMAD.lo", "IMAD.HI", "IMAD.WIDE(acc)", "IADD", "LOP3", "SHF", "SHL+ADD(LEA?)", "PRMT", "FFMA",
    "IMAD.WIDE+IADD", "IMAD.WIDE+IADD+LOP3", "IMAD.lo+IADD", "FFMA+IADD", "FFMA+IMAD.lo", "FFMA+IMAD.WIDE", "ADD.CC+ADDC", "ADD+ADD(IADD3?)",
    "2FFMA+IADD+LOP3", "IMAD.WIDE+4ALU", "MUL.WIDE+IADD"};
static const int NOPS[] = {1,1,1,1,1,1,1,1,1, 2,3,2,2,2,2,2,1, 4,5,2};

template <int KIND> void run(u32* d, int sms, int mhz) {
    const int iters = 1024, blocks = sms * 8;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<KIND><<<blocks, 256>>>(d, iters); cudaDeviceSynchronize();
    cudaEventRecord(a); k<KIND><<<blocks, 256>>>(d, iters); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double groups = (double)blocks * 256 * iters * 4 * 8;   // one "group" = one of each listed op
    double clk_per_group_warp = ms * 1e-3 * mhz * 1e6 * sms * 4 / (groups / 32);  // SMSP-cycles per warp-level group
    printf("%-22s %8.3f ms  %6.1f lane-groups/clk/SM  (%.2f SMSP-clk per warp-group of %d instr)\n", NAMES[KIND], ms,
           groups / (ms * 1e-3 * mhz * 1e6 * sms), clk_per_group_warp, NOPS[KIND]);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount, mhz = p.clockRate / 1000;
    printf("%s SMs=%d nominal %d MHz (rates assume the nominal clock)\n", p.name, sms, mhz);
    u32* d; cudaMalloc(&d, (size_t)sms * 8 * 256 * 4 + 4096); cudaMemset(d, 5, (size_t)sms * 8 * 256 * 4 + 4096);
    run<0>(d, sms, mhz); run<1>(d, sms, mhz); run<2>(d, sms, mhz); run<3>(d, sms, mhz); run<4>(d, sms, mhz);
    run<5>(d, sms, mhz); run<6>(d, sms, mhz); run<7>(d, sms, mhz); run<8>(d, sms, mhz); run<9>(d, sms, mhz);
    run<10>(d, sms, mhz); run<11>(d, sms, mhz); run<12>(d, sms, mhz); run<13>(d, sms, mhz); run<14>(d, sms, mhz);
    run<15>(d, sms, mhz); run<16>(d, sms, mhz); run<17>(d, sms, mhz); run<18>(d, sms, mhz); run<19>(d, sms, mhz);
    return 0;
}
