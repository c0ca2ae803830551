// gl.cuh -- Goldilocks field (p = 2^64 - 2^32 + 1) and its quadratic extension on the device.
//
// Replaces plonky2_field's GoldilocksField arithmetic (field/src/goldilocks_field.rs @ d2598bd, pinned
// by /root/reference/Cargo.lock:4912-4924) for the sm_100a kernels.  Integer
// pipe only: a 64x64->128 product is four IMAD.WIDE.U32, the reduction uses 2^64 == 2^32 - 1 and
// 2^96 == -1 (mod p).  "nc" variants return any representative in [0, 2^64); everything written to
// memory that the host can see is canonical (< p).
#pragma once
#include <cstdint>

namespace gl {

constexpr uint64_t P = 0xFFFFFFFF00000001ULL;
constexpr uint64_t EPS = 0xFFFFFFFFULL;
constexpr uint64_t GENERATOR = 7ULL;                        // MULTIPLICATIVE_GROUP_GENERATOR == coset_shift()
constexpr uint64_t POW2_GENERATOR = 1753635133440165772ULL; // order 2^32

__host__ __device__ __forceinline__ uint64_t canon(uint64_t x) { return x >= P ? x - P : x; }

// canonical + canonical -> canonical
__host__ __device__ __forceinline__ uint64_t add(uint64_t a, uint64_t b) {
    uint64_t s = a + b;
    return (s < a || s >= P) ? s + EPS : s; // s - p == s + eps (mod 2^64)
}
// any + canonical -> any (one operand canonical guarantees a single wrap)
__host__ __device__ __forceinline__ uint64_t add_nc(uint64_t a, uint64_t b_canonical) {
    uint64_t s = a + b_canonical;
    return s < a ? s + EPS : s;
}
// canonical - canonical -> canonical
__host__ __device__ __forceinline__ uint64_t sub(uint64_t a, uint64_t b) {
    uint64_t d = a - b;
    return a < b ? d - EPS : d; // d + p == d - eps (mod 2^64)
}
__host__ __device__ __forceinline__ uint64_t neg(uint64_t a) { return a ? P - a : 0; }

// (lo + 2^64 hi) mod p, any representative.  hi = hi_hi 2^32 + hi_lo:  lo - hi_hi + hi_lo * eps.
__host__ __device__ __forceinline__ uint64_t reduce128_nc(uint64_t lo, uint64_t hi) {
    uint64_t hi_hi = hi >> 32, hi_lo = hi & EPS;
    uint64_t t0 = lo - hi_hi;
    if (lo < hi_hi) t0 -= EPS;
    uint64_t t1 = (hi_lo << 32) - hi_lo; // hi_lo * eps
    uint64_t r = t0 + t1;
    if (r < t1) r += EPS;
    return r;
}

// ---- device-only canonical arithmetic on IADD3 carry chains (full-rate; no compare/select, which issue at half rate) ----
#ifdef __CUDACC__
// canonical - canonical -> canonical: a - b, and a borrow is repaid with + p == - eps (m = 0xffffffff = eps)
__device__ __forceinline__ uint64_t dsub(uint64_t a, uint64_t b) {
    const uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32);
    uint32_t r0, r1;
    asm("{\n\t"
        ".reg .u32 m;\n\t"
        "sub.cc.u32 %0,%2,%4;\n\t"
        "subc.cc.u32 %1,%3,%5;\n\t"
        "subc.u32 m,0,0;\n\t"
        "sub.cc.u32 %0,%0,m;\n\t"
        "subc.u32 %1,%1,0;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1));
    return ((uint64_t)r1 << 32) | r0;
}
// canonical + canonical -> canonical, as a - (p - b): p - b in (0, p] needs no correction of its own
__device__ __forceinline__ uint64_t dadd(uint64_t a, uint64_t b) {
    const uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32);
    uint32_t r0, r1;
    asm("{\n\t"
        ".reg .u32 m,n0,n1;\n\t"
        "sub.cc.u32 n0,1,%4;\n\t"
        "subc.u32 n1,0xffffffff,%5;\n\t"
        "sub.cc.u32 %0,%2,n0;\n\t"
        "subc.cc.u32 %1,%3,n1;\n\t"
        "subc.u32 m,0,0;\n\t"
        "sub.cc.u32 %0,%0,m;\n\t"
        "subc.u32 %1,%1,0;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(a0), "r"(a1), "r"(b0), "r"(b1));
    return ((uint64_t)r1 << 32) | r0;
}
// any representative -> canonical: x >= p exactly when x + eps carries out of 64 bits, and then x - p == x + eps
__device__ __forceinline__ uint64_t dcanon(uint64_t x) {
    const uint32_t x0 = (uint32_t)x, x1 = (uint32_t)(x >> 32);
    uint32_t r0, r1;
    asm("{\n\t"
        ".reg .u32 t,c;\n\t"
        "add.cc.u32 t,%2,0xffffffff;\n\t"
        "addc.cc.u32 t,%3,0;\n\t"
        "addc.u32 c,0,0;\n\t"
        "sub.u32 c,0,c;\n\t"
        "add.cc.u32 %0,%2,c;\n\t"
        "addc.u32 %1,%3,0;\n\t"
        "}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(x0), "r"(x1));
    return ((uint64_t)r1 << 32) | r0;
}
#endif

// (lo + 2^64 hi) mod p, any representative, for the device: with hi = 2^32 p3 + p2,
// V = lo + (p2 << 32) - p2 - p3 lies in (-2^33, 2^65); it is formed in 128-bit arithmetic (IADD3/IADD3.X
// chains, no multiplier), and its signed top word tw in {-1,0,1} weighs 2^64 == eps: one more 64-bit add
// that cannot wrap (after a carry the low 64 bits are < 2^64 - 2^33, after a borrow they are > 2^64 - 2^33).
__device__ __forceinline__ uint64_t reduce128_dev_nc(uint64_t lo, uint64_t hi) {
    const uint32_t p2 = (uint32_t)hi, p3 = (uint32_t)(hi >> 32);
    const __int128 V = (__int128)lo + ((__int128)(uint64_t)p2 << 32) - (__int128)(uint64_t)p2 - (__int128)(uint64_t)p3;
    const int tw = (int)(V >> 64);
    const uint64_t adj = ((uint64_t)(uint32_t)(tw >> 31) << 32) | (uint32_t)(-tw); // tw * eps mod 2^64
    return (uint64_t)V + adj;
}
// four IMAD.WIDE.U32 (the multiplier issues one per 4 clocks per sub-partition: they bound the S-box)
__device__ __forceinline__ uint64_t mul_nc(uint64_t a, uint64_t b) {
    const unsigned __int128 p = (unsigned __int128)a * b;
    return reduce128_dev_nc((uint64_t)p, (uint64_t)(p >> 64));
}
__device__ __forceinline__ uint64_t mul(uint64_t a, uint64_t b) { return dcanon(mul_nc(a, b)); }
// a * c for a 32-bit c (small constants: MDS entries, limb weights): two IMAD.WIDE; the 96-bit product lo + 2^64 h has
// h < 2^32, and 2^64 == eps, so it is lo + h * eps with one conditional repair -- half the work of a full product
__device__ __forceinline__ uint64_t mul_u32(uint64_t a, uint32_t c) {
    const uint64_t p0 = (uint64_t)(uint32_t)a * c, p1 = (a >> 32) * c;
    const uint64_t mid = p1 + (p0 >> 32);                       // < 2^64: (2^32-1)^2 + 2^32 - 1
    const uint64_t lo = (mid << 32) | (uint32_t)p0, h = mid >> 32; // a * c = lo + 2^64 h
    const uint64_t t = (h << 32) - h;                            // h * eps < 2^64
    uint64_t r = lo + t;
    if (r < t) r += EPS;                                         // wrapped once: + 2^64 == + eps, cannot wrap again
    return dcanon(r);
}

// x * 2^s (mod p) for a compile-time s in [0, 96), canonical in, canonical out, no multiplier: 2 generates the
// 192-element subgroup of F_p^* (2^96 == -1), so every root of unity of order <= 64 is a power of two and the
// twiddles inside a radix-2^R register step are shifts.  z = x << (s % 32) is three words; the word offset
// s / 32 is folded with 2^64 == 2^32 - 1, 2^96 == -1, 2^128 == -2^32 into a signed 128-bit sum whose top word
// tw in {-1,0,1} is closed exactly as in reduce128_dev_nc.
__device__ __forceinline__ uint64_t mul_pow2(uint64_t x, const int s) {
    if (s == 0) return x;
    const int q = s >> 5, r = s & 31;
    const uint32_t x0 = (uint32_t)x, x1 = (uint32_t)(x >> 32);
    const uint64_t z0 = r ? (uint32_t)(x0 << r) : x0;
    const uint64_t z1 = r ? __funnelshift_l(x0, x1, r) : x1;
    const uint64_t z2 = r ? (x1 >> (32 - r)) : 0u;
    __int128 V;
    if (q == 0)
        V = (__int128)(z0 | (z1 << 32)) + (__int128)(z2 << 32) - (__int128)z2;
    else if (q == 1)
        V = ((__int128)(z0 + z1) << 32) - (__int128)z1 - (__int128)z2;
    else
        V = (((__int128)z0 - (__int128)z2) << 32) - (__int128)z0 - (__int128)z1;
    const int tw = (int)(V >> 64);
    const uint64_t adj = ((uint64_t)(uint32_t)(tw >> 31) << 32) | (uint32_t)(-tw);
    return dcanon((uint64_t)V + adj);
}
__device__ __forceinline__ uint64_t sqr_nc(uint64_t a) { return mul_nc(a, a); }

__device__ __forceinline__ uint64_t pow(uint64_t b, uint64_t e) {
    uint64_t r = 1;
    while (e) {
        if (e & 1) r = mul(r, b);
        b = mul(b, b);
        e >>= 1;
    }
    return r;
}
__device__ __forceinline__ uint64_t inv(uint64_t a) { return pow(a, P - 2); }

// ---- host-side scalar helpers (setup only: roots, shifts, table seeds; never on a data path) ----
inline uint64_t h_mul(uint64_t a, uint64_t b) {
    unsigned __int128 m = (unsigned __int128)a * b;
    return canon(reduce128_nc((uint64_t)m, (uint64_t)(m >> 64)));
}
inline uint64_t h_pow(uint64_t b, uint64_t e) {
    uint64_t r = 1;
    while (e) {
        if (e & 1) r = h_mul(r, b);
        b = h_mul(b, b);
        e >>= 1;
    }
    return r;
}
inline uint64_t h_inv(uint64_t a) { return h_pow(a, P - 2); }
inline uint64_t h_root_of_unity(int n_log) { // primitive_root_of_unity(n_log)
    uint64_t r = POW2_GENERATOR;
    for (int i = n_log; i < 32; i++) r = h_mul(r, r);
    return r;
}

// ---- quadratic extension F_p[X]/(X^2 - 7), element (a0, a1) ----
struct Ext2 {
    uint64_t a0, a1;
};
__device__ __forceinline__ Ext2 ext_add(Ext2 x, Ext2 y) { return {add(x.a0, y.a0), add(x.a1, y.a1)}; }
__device__ __forceinline__ Ext2 ext_sub(Ext2 x, Ext2 y) { return {sub(x.a0, y.a0), sub(x.a1, y.a1)}; }
__device__ __forceinline__ Ext2 ext_mul(Ext2 x, Ext2 y) {
    uint64_t c0 = add(mul(x.a0, y.a0), mul(7, mul(x.a1, y.a1)));
    uint64_t c1 = add(mul(x.a0, y.a1), mul(x.a1, y.a0));
    return {c0, c1};
}
__device__ __forceinline__ Ext2 ext_scalar(Ext2 x, uint64_t s) { return {mul(x.a0, s), mul(x.a1, s)}; }
__device__ __forceinline__ Ext2 ext_pow(Ext2 b, uint64_t e) {
    Ext2 r = {1, 0};
    while (e) {
        if (e & 1) r = ext_mul(r, b);
        b = ext_mul(b, b);
        e >>= 1;
    }
    return r;
}

} // namespace gl
