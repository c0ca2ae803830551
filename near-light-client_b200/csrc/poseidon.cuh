// poseidon.cuh -- width-12 Poseidon permutation over Goldilocks, one permutation per thread.
//
// Replaces plonky2's PoseidonPermutation / PoseidonHash (plonky2/src/hash/poseidon.rs,
// poseidon_goldilocks.rs, hashing.rs @ d2598bd), the hasher bound at
// /root/reference/vendor/starkyx/starkyx/src/plonky2/stark/config.rs:24,104.
//
// Integer-pipe kernel body: the state lives in 24 registers, values stay non-canonical ("nc") between
// layers and are canonicalised once at the end.  The MDS layer exploits the tiny circulant constants:
// each lane is split into 32-bit halves and the two 12-term dot products are accumulated in plain u64
// (12 * 2^32 * 41 < 2^42) with IMAD.WIDE.U32, then recombined with a single cheap reduction.
#pragma once
#include "gl.cuh"
#include "poseidon_constants.h"

namespace poseidon {

constexpr int WIDTH = 12;
constexpr int RATE = 8;
constexpr int HALF_FULL = 4;
constexpr int N_PARTIAL = 22;

static __constant__ uint64_t c_rc[NLC_POSEIDON_N_RC] = NLC_POSEIDON_RC_INIT;

__device__ __forceinline__ uint64_t sbox7(uint64_t x) {
    uint64_t x2 = gl::sqr_nc(x), x4 = gl::sqr_nc(x2), x3 = gl::mul_nc(x, x2);
    return gl::mul_nc(x3, x4);
}

// out[r] = sum_i s[(i+r)%12] * CIRC[i] + s[r] * DIAG[r]  (+ rc[r] when ADD_RC: the NEXT round's constant
// layer is folded into the accumulators, where a 64-bit add cannot overflow).
template <bool ADD_RC>
__device__ __forceinline__ void mds_layer(uint64_t (&s)[WIDTH], const uint64_t* __restrict__ rc) {
    constexpr uint32_t CIRC[WIDTH] = {17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20};
    uint32_t lo[WIDTH], hi[WIDTH];
#pragma unroll
    for (int i = 0; i < WIDTH; i++) {
        lo[i] = (uint32_t)s[i];
        hi[i] = (uint32_t)(s[i] >> 32);
    }
#pragma unroll
    for (int r = 0; r < WIDTH; r++) {
        uint64_t alo = 0, ahi = 0;
        if (ADD_RC) {
            const uint64_t c = rc[r];
            alo = (uint32_t)c;
            ahi = c >> 32;
        }
#pragma unroll
        for (int i = 0; i < WIDTH; i++) {
            alo += (uint64_t)lo[(i + r) % WIDTH] * CIRC[i];
            ahi += (uint64_t)hi[(i + r) % WIDTH] * CIRC[i];
        }
        if (r == 0) { // MDS_MATRIX_DIAG = [8, 0, ...]
            alo += (uint64_t)lo[0] * 8u;
            ahi += (uint64_t)hi[0] * 8u;
        }
        // value = alo + 2^32 ahi with alo, ahi < 2^43.  The top 11 bits of ahi weigh 2^64 == eps and go into alo
        // (no overflow); the low word of ahi is added to the high word of that sum, and a carry out of it is
        // one more 2^64 == eps, added to the (now small) wrapped value.
        const uint32_t ahl = (uint32_t)ahi, ahh = (uint32_t)(ahi >> 32);
        const uint64_t a2 = alo + (uint64_t)ahh * 0xffffffffu;
        const uint32_t a2l = (uint32_t)a2, a2h = (uint32_t)(a2 >> 32);
        uint32_t r0, r1;
        asm("{\n\t"
            ".reg .u32 h,m;\n\t"
            "add.cc.u32 h,%3,%4;\n\t"
            "addc.u32 m,0,0;\n\t"
            "sub.u32 m,0,m;\n\t" // m = carry ? 0xffffffff (= eps) : 0
            "add.cc.u32 %0,%2,m;\n\t"
            "addc.u32 %1,h,0;\n\t"
            "}"
            : "=r"(r0), "=r"(r1)
            : "r"(a2l), "r"(a2h), "r"(ahl));
        s[r] = ((uint64_t)r1 << 32) | r0;
    }
}

// The permutation in plonky2's naive form (constant layer, S-box layer, MDS layer per round), with every
// constant layer but the first folded into the preceding MDS layer.
__device__ __forceinline__ void permute_nc(uint64_t (&s)[WIDTH]) {
#pragma unroll
    for (int i = 0; i < WIDTH; i++) s[i] = gl::add_nc(s[i], c_rc[i]);
    const uint64_t* rc = c_rc + WIDTH;
#pragma unroll 1
    for (int r = 0; r < HALF_FULL; r++) {
#pragma unroll
        for (int i = 0; i < WIDTH; i++) s[i] = sbox7(s[i]);
        mds_layer<true>(s, rc);
        rc += WIDTH;
    }
#pragma unroll 1
    for (int r = 0; r < N_PARTIAL; r++) {
        s[0] = sbox7(s[0]);
        mds_layer<true>(s, rc);
        rc += WIDTH;
    }
#pragma unroll 1
    for (int r = 0; r < HALF_FULL - 1; r++) {
#pragma unroll
        for (int i = 0; i < WIDTH; i++) s[i] = sbox7(s[i]);
        mds_layer<true>(s, rc);
        rc += WIDTH;
    }
#pragma unroll
    for (int i = 0; i < WIDTH; i++) s[i] = sbox7(s[i]);
    mds_layer<false>(s, nullptr);
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
