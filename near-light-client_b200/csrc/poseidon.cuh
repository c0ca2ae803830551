// poseidon.cuh -- width-12 Poseidon permutation over Goldilocks, one permutation per thread.
//
// Replaces plonky2's PoseidonPermutation / PoseidonHash (plonky2/src/hash/poseidon.rs,
// poseidon_goldilocks.rs, hashing.rs @ d2598bd), the hasher bound at
// /root/reference/vendor/starkyx/starkyx/src/plonky2/stark/config.rs:24,104.
//
// Instruction budget (measured on B200, tools/ubench/ubench_pipes.cu): IMAD.WIDE/IMAD.HI issue once per 4
// clocks per SM sub-partition, 32-bit IMAD/LOP3/SHF/PRMT once per 2, IADD3 once per clock.  So the S-box
// keeps its 64x64 products on IMAD.WIDE (nothing cheaper exists), but the MDS layer uses NO multiplier:
// the circulant was chosen by plonky2 so that its two-level even/odd split (y^12-1 = (y^6-1)(y^6+1),
// y^6-1 = (y^3-1)(y^3+1)) has only powers of two as constants.  Each lane is cut into three limbs of
// 22/21/21 bits, the 12-point circular correlation runs on the limbs in plain 32-bit wrap-around
// arithmetic (adds and small shifts; the true result 264 * 2^22 < 2^32 fits), and the three limb sums are
// glued back with three IMAD.WIDE and one fold of the bits above 2^64.  The next round's constants ride
// along in limb form.  Values stay non-canonical ("nc", any representative below 2^64) between layers.
#pragma once
#include "gl.cuh"
#include "poseidon_constants.h"

// Pipe balance knob (tools/ubench/ubench_perm.cu): bit k set => the adds of network level k get a runtime-zero third
// addend, which forces IADD3 (ALU pipe) where ptxas would otherwise pick IMAD.IADD (multiplier pipe).  With the
// carry-chain recombination the multiplier pipe has slack and ptxas' own split wins: 0 measured best (1.17 vs 1.10 G/s).
#ifndef NLC_ZMASK
#define NLC_ZMASK 0
#endif

namespace poseidon {

constexpr int WIDTH = 12;
constexpr int RATE = 8;
constexpr int HALF_FULL = 4;
constexpr int N_PARTIAL = 22;
constexpr int N_ROUNDS = 2 * HALF_FULL + N_PARTIAL;

static __constant__ uint64_t c_rc[NLC_POSEIDON_N_RC] = NLC_POSEIDON_RC_INIT;

// Every limb-form constant carries a bias (BIAS0, BIAS1, BIAS2) with BIAS0 + 2^22 BIAS1 + 2^43 BIAS2 = 256 p: it
// changes no value mod p, and it keeps every limb sum of a layer strictly positive even when a loosely normalised
// input limb (partial rounds, below) has dipped under zero -- so carries are plain logical shifts and glue3's
// unsigned preconditions (O0, O1 < 2^32, O2 < 2^31) hold with room to spare.
constexpr uint32_t BIAS0 = (1u << 22) + 256u, BIAS1 = 7u * (1u << 18) - 1u, BIAS2 = (1u << 29) - 1u;
static_assert(((unsigned __int128)BIAS0 + ((unsigned __int128)BIAS1 << 22) + ((unsigned __int128)BIAS2 << 43)) %
                      0xFFFFFFFF00000001ULL == 0, "limb bias must be a multiple of p");
// Round constants in limb form: v[(r*3 + limb)*12 + lane]; one extra all-zero round closes the last layer.
struct RcLimbs {
    uint32_t v[(N_ROUNDS + 1) * 3 * WIDTH];
};
constexpr RcLimbs make_rc_limbs() {
    RcLimbs t{};
    constexpr uint64_t rc[NLC_POSEIDON_N_RC] = NLC_POSEIDON_RC_INIT;
    for (int r = 0; r < N_ROUNDS; r++)
        for (int i = 0; i < WIDTH; i++) {
            const uint64_t c = rc[r * WIDTH + i];
            t.v[(r * 3 + 0) * WIDTH + i] = (uint32_t)(c & 0x3FFFFFu) + BIAS0;
            t.v[(r * 3 + 1) * WIDTH + i] = (uint32_t)((c >> 22) & 0x1FFFFFu) + BIAS1;
            t.v[(r * 3 + 2) * WIDTH + i] = (uint32_t)(c >> 43) + BIAS2;
        }
    for (int i = 0; i < WIDTH; i++) { // the closing all-zero round carries the bias too
        t.v[(N_ROUNDS * 3 + 0) * WIDTH + i] = BIAS0;
        t.v[(N_ROUNDS * 3 + 1) * WIDTH + i] = BIAS1;
        t.v[(N_ROUNDS * 3 + 2) * WIDTH + i] = BIAS2;
    }
    return t;
}
static __constant__ RcLimbs c_rc_limbs = make_rc_limbs();
// Always 0, but not a compile-time constant: "a + b + z" has three live inputs, so ptxas must emit IADD3
// (ALU pipe, one per clock) and cannot pick IMAD.IADD (the multiplier pipe, which the S-boxes need).
static __constant__ uint32_t c_zero;

__device__ __forceinline__ uint64_t sbox7(uint64_t x) {
    uint64_t x2 = gl::sqr_nc(x), x4 = gl::sqr_nc(x2), x3 = gl::mul_nc(x, x2);
    return gl::mul_nc(x3, x4);
}

// o[r] = sum_i x[(i+r)%12] * CIRC[i] + [r==0] 8 x[0] + rc[r]  (mod 2^32), CIRC = 17 15 41 16 2 28 13 13 39 18 34 20.
// Written as the cyclic convolution with k[m] = CIRC[-m]; even/odd split twice, halvings folded into the
// constants:  k+ + = [16,32,16] (cyclic 3),  k+ - = [-1,-8,2] (negacyclic 3),  k- = [2,-4,16,1,-1,-1] (negacyclic 6).
__device__ __forceinline__ void circ12(const uint32_t (&x)[WIDTH], uint32_t (&o)[WIDTH], const uint32_t* __restrict__ rc, const uint32_t z) {
    uint32_t p[6], m[6];
#pragma unroll
    for (int j = 0; j < 6; j++) {
        p[j] = x[j] + x[j + 6] + (NLC_ZMASK & 1 ? z : 0u);
        m[j] = x[j] - x[j + 6] + (NLC_ZMASK & 2 ? z : 0u);
    }
    uint32_t t[3], w[3];
#pragma unroll
    for (int j = 0; j < 3; j++) {
        t[j] = p[j] + p[j + 3] + (NLC_ZMASK & 4 ? z : 0u);
        w[j] = p[j] - p[j + 3] + (NLC_ZMASK & 4 ? z : 0u);
    }
    const uint32_t S = t[0] + t[1] + t[2];
    const uint32_t ap[3] = {S + t[2] + (NLC_ZMASK & 8 ? z : 0u), S + t[0] + (NLC_ZMASK & 8 ? z : 0u), S + t[1] + (NLC_ZMASK & 8 ? z : 0u)}; // x16
    const uint32_t am[3] = {8u * w[2] - 2u * w[1] - w[0], 0u - 8u * w[0] - 2u * w[2] - w[1],
                            2u * w[0] - 8u * w[1] - w[2]};
    uint32_t A[6];
#pragma unroll
    for (int j = 0; j < 3; j++) {
        A[j] = 16u * ap[j] + am[j];
        A[j + 3] = 16u * ap[j] - am[j];
    }
    uint32_t B[6];
    B[0] = 2u * m[0] + 4u * m[5] - 16u * m[4] - m[3] + m[2] + m[1];
    B[1] = 2u * m[1] - 4u * m[0] - 16u * m[5] - m[4] + m[3] + m[2];
    B[2] = 2u * m[2] - 4u * m[1] + 16u * m[0] - m[5] + m[4] + m[3];
    B[3] = 2u * m[3] - 4u * m[2] + 16u * m[1] + m[0] + m[5] + m[4];
    B[4] = 2u * m[4] - 4u * m[3] + 16u * m[2] + m[1] - m[0] + m[5];
    B[5] = 2u * m[5] - 4u * m[4] + 16u * m[3] + m[2] - m[1] - m[0];
#pragma unroll
    for (int j = 0; j < 6; j++) {
        o[j] = A[j] + B[j] + rc[j];
        o[j + 6] = A[j] - B[j] + rc[j + 6];
    }
    o[0] += 8u * x[0]; // MDS_MATRIX_DIAG = [8, 0, ...]
}

// V = O0 + 2^22 O1 + 2^43 O2 (mod p), any representative.  With O2 = 2^21 h + l the top part h 2^64 == h 2^32 - h is
// folded BEFORE the limbs are assembled: V == (O0 - h) + 2^22 O1 + 2^32 h + 2^43 l, where O0 - h >= 0 because every
// limb sum carries BIAS0 > 2^22 > h, and the sum is below 2^64 + 2^53: at most one carry out of 64 bits, repaid with
// + eps (2^64 == eps), which cannot wrap again.  12 instructions.
__device__ __forceinline__ uint64_t glue3(uint32_t O0, uint32_t O1, uint32_t O2) {
    uint32_t r0, r1;
    asm("{\n\t"
        ".reg .u32 h,a,b,c,t,w0,w1,k,m;\n\t"
        "shr.u32 h,%4,21;\n\t"
        "shl.b32 a,%3,22;\n\t"
        "shr.u32 b,%3,10;\n\t"
        "shl.b32 c,%4,11;\n\t"
        "sub.u32 t,%2,h;\n\t"
        "add.u32 b,b,h;\n\t"
        "add.cc.u32 w0,t,a;\n\t"
        "addc.cc.u32 w1,b,c;\n\t"
        "addc.u32 k,0,0;\n\t"
        "sub.u32 m,0,k;\n\t"
        "add.cc.u32 %0,w0,m;\n\t"
        "addc.u32 %1,w1,0;\n\t"
        "}"
        : "=r"(r0), "=r"(r1)
        : "r"(O0), "r"(O1), "r"(O2));
    return ((uint64_t)r1 << 32) | r0;
}

__device__ __forceinline__ void split3(uint64_t v, uint32_t& l0, uint32_t& l1, uint32_t& l2) {
    l0 = (uint32_t)v & 0x3FFFFFu;
    l1 = (uint32_t)(v >> 22) & 0x1FFFFFu;
    l2 = (uint32_t)(v >> 43);
}

// ---- the MDS layer of the FULL rounds on the FP64 pipe: built, bit-exact, measured, NOT faster -> off by default -------
// B200's FP64 pipe issues DFMA / DADD at the rate of the integer pipes (tools/ubench/ubench_fp64.cu: 63 per clock per
// SM, and it co-issues with IMAD: 119 per clock for the two streams together) and the hash leaves it idle.  Integers
// below 2^53 are exact in a double, so a lane is cut into its two 32-bit halves, the 12-point circular correlation runs
// on each half in doubles (inputs < 2^32, every intermediate < 2^41), and the next round's constants ride along with a
// bias that makes the result's mantissa the integer itself: 2 planes instead of 3, no masks, a cheaper glue.  SASS of a
// full round: 1325 -> 1264 instructions (204 of them DADD / DFMA), busiest integer pipe 1546 -> 1244 clocks -- and the
// permutation rate does not move (1.330 -> 1.332 G/s; 6.61 -> 6.80 ms inside the commit): the kernel is bound by
// issue-slot utilisation (~0.73 instructions per clock per sub-partition), not by any one pipe, so only fewer
// instructions would help, and 61 of 20 100 is noise.  Kept as a compile-time variant (-DNLC_MDS_FP64=1).
#ifndef NLC_MDS_FP64
#define NLC_MDS_FP64 0
#endif

// J0 = O0 + 2^51 + 2^19, J1 = O1 + 2^51 - 2^20 are what the biased planes hold in their mantissas; the biases cancel:
// 2^51 + 2^19 + 2^32 (2^51 - 2^20) = 2^83 + 2^51 - 2^52 + 2^19 == (2^51 - 2^19) + 2^51 - 2^52 + 2^19 = 0  (2^64 == 2^32 - 1).
constexpr double FP64_BIAS0 = 6755399441055744.0 + 524288.0;  // 1.5 * 2^52 + 2^19
constexpr double FP64_BIAS1 = 6755399441055744.0 - 1048576.0; // 1.5 * 2^52 - 2^20
struct RcDoubles {
    double v[(N_ROUNDS + 1) * 2 * WIDTH]; // v[(r*2 + plane)*12 + lane], biased; round N_ROUNDS = the closing zero round
};
constexpr RcDoubles make_rc_doubles() {
    RcDoubles t{};
    constexpr uint64_t rc[NLC_POSEIDON_N_RC] = NLC_POSEIDON_RC_INIT;
    for (int r = 0; r <= N_ROUNDS; r++)
        for (int i = 0; i < WIDTH; i++) {
            const uint64_t c = r < N_ROUNDS ? rc[r * WIDTH + i] : 0;
            t.v[(r * 2 + 0) * WIDTH + i] = (double)(uint32_t)c + FP64_BIAS0;
            t.v[(r * 2 + 1) * WIDTH + i] = (double)(uint32_t)(c >> 32) + FP64_BIAS1;
        }
    return t;
}
static __constant__ RcDoubles c_rc_doubles = make_rc_doubles();

__device__ __forceinline__ double u32_to_double(uint32_t v) { // exact: 2^52 + v has v in its low mantissa word
    return __hiloint2double(0x43300000, (int)v) - 4503599627370496.0;
}

// the same two-level even/odd split as circ12, in doubles (fma for the power-of-two weights)
__device__ __forceinline__ void circ12d(const double (&x)[WIDTH], double (&o)[WIDTH], const double* __restrict__ rc) {
    double p[6], m[6];
#pragma unroll
    for (int j = 0; j < 6; j++) {
        p[j] = x[j] + x[j + 6];
        m[j] = x[j] - x[j + 6];
    }
    double t[3], w[3];
#pragma unroll
    for (int j = 0; j < 3; j++) {
        t[j] = p[j] + p[j + 3];
        w[j] = p[j] - p[j + 3];
    }
    const double S = t[0] + t[1] + t[2];
    const double ap[3] = {S + t[2], S + t[0], S + t[1]};
    const double am[3] = {fma(8.0, w[2], fma(-2.0, w[1], -w[0])), fma(-8.0, w[0], fma(-2.0, w[2], -w[1])),
                          fma(2.0, w[0], fma(-8.0, w[1], -w[2]))};
    double A[6];
#pragma unroll
    for (int j = 0; j < 3; j++) {
        A[j] = fma(16.0, ap[j], am[j]);
        A[j + 3] = fma(16.0, ap[j], -am[j]);
    }
    double B[6];
    B[0] = fma(-16.0, m[4], fma(4.0, m[5], fma(2.0, m[0], (m[2] + m[1]) - m[3])));
    B[1] = fma(-16.0, m[5], fma(-4.0, m[0], fma(2.0, m[1], (m[3] + m[2]) - m[4])));
    B[2] = fma(16.0, m[0], fma(-4.0, m[1], fma(2.0, m[2], (m[4] + m[3]) - m[5])));
    B[3] = fma(16.0, m[1], fma(-4.0, m[2], fma(2.0, m[3], (m[0] + m[5]) + m[4])));
    B[4] = fma(16.0, m[2], fma(-4.0, m[3], fma(2.0, m[4], (m[1] - m[0]) + m[5])));
    B[5] = fma(16.0, m[3], fma(-4.0, m[4], fma(2.0, m[5], (m[2] - m[1]) - m[0])));
#pragma unroll
    for (int j = 0; j < 6; j++) {
        o[j] = (A[j] + rc[j]) + B[j];
        o[j + 6] = (A[j] + rc[j + 6]) - B[j];
    }
    o[0] = fma(8.0, x[0], o[0]); // MDS_MATRIX_DIAG = [8, 0, ...]
}

// (J0, J1) -> J0 + 2^32 J1 mod p, any representative.  J = hi 2^32 + lo with hi < 2^20 (the mantissa's top 20 bits):
// 2^32 J1 = 2^32 lo1 + 2^64 hi1 == 2^32 (lo1 + hi1) - hi1; J0 >= 2^50 so J0 - hi1 cannot borrow, and adding
// lo1 + hi1 < 2^33 to a high word < 2^20 carries out at most once, repaid with + eps (which cannot wrap again).
__device__ __forceinline__ uint64_t glue2(double J0, double J1) {
    const uint32_t lo0 = (uint32_t)__double2loint(J0), hi0 = (uint32_t)__double2hiint(J0) & 0xFFFFFu;
    const uint32_t lo1 = (uint32_t)__double2loint(J1), hi1 = (uint32_t)__double2hiint(J1) & 0xFFFFFu;
    uint32_t r0, r1;
    asm("{\n\t"
        ".reg .u32 t0,t1,k,m;\n\t"
        "sub.cc.u32 t0,%2,%5;\n\t"
        "subc.u32 t1,%3,0;\n\t"
        "add.u32 t1,t1,%5;\n\t"
        "add.cc.u32 t1,t1,%4;\n\t"
        "addc.u32 k,0,0;\n\t"
        "sub.u32 m,0,k;\n\t"
        "add.cc.u32 %0,t0,m;\n\t"
        "addc.u32 %1,t1,0;\n\t"
        "}"
        : "=r"(r0), "=r"(r1)
        : "r"(lo0), "r"(hi0), "r"(lo1), "r"(hi1));
    return ((uint64_t)r1 << 32) | r0;
}

// s <- MDS s + rc of round `next_round`, on the FP64 pipe (full rounds)
__device__ __forceinline__ void mds_layer_fp64(uint64_t (&s)[WIDTH], const int next_round) {
    double L0[WIDTH], L1[WIDTH], O0[WIDTH], O1[WIDTH];
#pragma unroll
    for (int i = 0; i < WIDTH; i++) {
        L0[i] = u32_to_double((uint32_t)s[i]);
        L1[i] = u32_to_double((uint32_t)(s[i] >> 32));
    }
    const double* rc = c_rc_doubles.v + next_round * 2 * WIDTH;
    circ12d(L0, O0, rc);
    circ12d(L1, O1, rc + WIDTH);
#pragma unroll
    for (int i = 0; i < WIDTH; i++) s[i] = glue2(O0[i], O1[i]);
}

// s <- MDS s + rc, rc = 36 limb words of the NEXT round's constants (full rounds: 64-bit lanes in and out).
__device__ __forceinline__ void mds_layer(uint64_t (&s)[WIDTH], const uint32_t* __restrict__ rc, const uint32_t z) {
    uint32_t L0[WIDTH], L1[WIDTH], L2[WIDTH], O0[WIDTH], O1[WIDTH], O2[WIDTH];
#pragma unroll
    for (int i = 0; i < WIDTH; i++) split3(s[i], L0[i], L1[i], L2[i]);
    circ12(L0, O0, rc, z);
    circ12(L1, O1, rc + WIDTH, z);
    circ12(L2, O2, rc + 2 * WIDTH, z);
#pragma unroll
    for (int i = 0; i < WIDTH; i++) s[i] = glue3(O0[i], O1[i], O2[i]);
}

// plonky2's naive round structure (constant layer, S-box layer, MDS layer); every constant layer but the
// first is folded into the preceding MDS layer.
//
// Full rounds take and return 64-bit lanes (every lane meets an S-box).  In the 22 partial rounds only lane 0
// does: lanes 1..11 go from one MDS layer straight into the next, so they STAY in limb form.  After a layer
// the limb sums O_k (< 2^31 in magnitude) are brought back under ~2^22.6 by one parallel carry step, no ripple:
//   L0 = (O0 & m22) - c2,  L1 = (O1 & m21) + c0 + (c2 << 10),  L2 = (O2 & m21) + c1,   c_k = O_k >> {22,21,21}
// (the carry out of the top limb weighs 2^64 == 2^32 - 1: +2^10 units of limb 1, -1 unit of limb 0; L0 may dip a
// few hundred below zero, which the bias of the next layer's constants absorbs).  10 instructions per lane instead of glue (16) + split (5);
// only lane 0 is glued for its S-box.  264 * 2^22.6 + 2^22 < 2^31, so the next layer cannot overflow.
// One body per round kind, both visited from one outer loop, keeps the code inside the instruction cache.
__device__ __forceinline__ void permute_nc(uint64_t (&s)[WIDTH]) {
#pragma unroll
    for (int i = 0; i < WIDTH; i++) s[i] = gl::add_nc(s[i], c_rc[i]);
    const uint32_t* rc = c_rc_limbs.v;
    const uint32_t z = c_zero;
    int rnd = 0; // rounds done: the layer of round r adds the constants of round r + 1
#pragma unroll 1
    for (int phase = 0; phase < 2; phase++) {
#pragma unroll 1
        for (int r = 0; r < HALF_FULL; r++) {
#pragma unroll
            for (int i = 0; i < WIDTH; i++) s[i] = sbox7(s[i]);
            rc += 3 * WIDTH;
            rnd++;
#if NLC_MDS_FP64
            mds_layer_fp64(s, rnd);
#else
            mds_layer(s, rc, z);
#endif
        }
        if (phase) break;
        rnd += N_PARTIAL;
        uint32_t L0[WIDTH], L1[WIDTH], L2[WIDTH];
#pragma unroll
        for (int i = 1; i < WIDTH; i++) split3(s[i], L0[i], L1[i], L2[i]);
        uint64_t s0 = s[0];
        uint32_t O0[WIDTH], O1[WIDTH], O2[WIDTH];
#pragma unroll 1
        for (int r = 0;; r++) {
            s0 = sbox7(s0);
            split3(s0, L0[0], L1[0], L2[0]);
            rc += 3 * WIDTH;
            circ12(L0, O0, rc, z);
            circ12(L1, O1, rc + WIDTH, z);
            circ12(L2, O2, rc + 2 * WIDTH, z);
            if (r == N_PARTIAL - 1) break; // the last layer's sums go back to 64-bit lanes below
            s0 = glue3(O0[0], O1[0], O2[0]);
#pragma unroll
            for (int i = 1; i < WIDTH; i++) {
                const uint32_t c0 = O0[i] >> 22, c1 = O1[i] >> 21, c2 = O2[i] >> 21; // O_k > 0 thanks to the bias
                L0[i] = (O0[i] & 0x3FFFFFu) - c2;
                L1[i] = (O1[i] & 0x1FFFFFu) + c0 + (c2 << 10);
                L2[i] = (O2[i] & 0x1FFFFFu) + c1;
            }
        }
#pragma unroll
        for (int i = 0; i < WIDTH; i++) s[i] = glue3(O0[i], O1[i], O2[i]);
    }
}

// canonical output
__device__ __forceinline__ void permute(uint64_t (&s)[WIDTH]) {
    permute_nc(s);
#pragma unroll
    for (int i = 0; i < WIDTH; i++) s[i] = gl::canon(s[i]);
}

// compress / two_to_one: state = [l0..l3, r0..r3, 0, 0, 0, 0]
__device__ __forceinline__ void two_to_one(const uint64_t* l, const uint64_t* r, uint64_t* out) {
    uint64_t s[WIDTH];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        s[i] = l[i];
        s[4 + i] = r[i];
        s[8 + i] = 0;
    }
    permute_nc(s);
#pragma unroll
    for (int i = 0; i < 4; i++) out[i] = gl::canon(s[i]);
}

} // namespace poseidon
