// transcript.cuh -- device-resident Fiat-Shamir transcript (plonky2::iop::challenger::Challenger).
//
// The sponge chain is inherently sequential, so the latency of ONE permutation is what matters: a
// single warp holds the state one lane per element (lanes 0..11), S-boxes run in parallel across lanes
// and the MDS layer is 12 shuffles.  Keeping the transcript on the device lets prove_openings enqueue
// commit rounds, folds, proof-of-work and query gathers without a single host round trip.
// Semantics: SURVEY.md Appendix A.8; order of use /root/reference/vendor/starkyx/starkyx/src/plonky2/
// stark/prover.rs:67,94-100,181-183,193.
#pragma once
#include "common.cuh"
#include "poseidon.cuh"

namespace transcript {

// One permutation per half-warp: lanes base..base+11 (base = lane & 16) pass s_i and receive the permuted s_i
// (canonical); lanes 12..15 of each half idle along.  All 32 lanes must call.  The latency of ONE permutation is
// what this form is for (transcript, top of a Merkle tree): S-boxes run across lanes, the MDS layer is 12 shuffles.
__device__ __forceinline__ u64 warp_permute(u64 x, int lane) {
    constexpr u32 CIRC[12] = {17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20};
    const int base = lane & 16, l16 = lane & 15;
    const int l12 = l16 < 12 ? l16 : 0;
#pragma unroll 1
    for (int r = 0; r < 30; r++) {
        x = gl::add_nc(x, poseidon::c_rc[r * 12 + l12]);
        const bool full = r < 4 || r >= 26;
        if (full || l16 == 0) x = poseidon::sbox7(x);
        const u32 lo = (u32)x, hi = (u32)(x >> 32);
        u64 alo = 0, ahi = 0;
#pragma unroll
        for (int i = 0; i < 12; i++) {
            int src = l12 + i;
            src = base + (src >= 12 ? src - 12 : src);
            u32 l = __shfl_sync(0xffffffffu, lo, src);
            u32 h = __shfl_sync(0xffffffffu, hi, src);
            alo += (u64)l * CIRC[i];
            ahi += (u64)h * CIRC[i];
        }
        if (l16 == 0) {
            alo += (u64)lo * 8u;
            ahi += (u64)hi * 8u;
        }
        u64 l = alo + (ahi << 32);
        u64 h = (ahi >> 32) + (l < alo ? 1 : 0);
        u64 t1 = (h << 32) - h;
        u64 v = l + t1;
        if (v < t1) v += gl::EPS;
        x = v;
    }
    return gl::canon(x);
}

// Challenger::duplexing, executed by one full warp on a challenger in global/shared memory.
__device__ __forceinline__ void duplex(nlc_challenger* c, int lane) {
    u64 x = 0;
    if (lane < 12) {
        x = c->sponge_state[lane];
        if (lane < 8 && (u32)lane < c->input_len) x = c->input_buffer[lane]; // overwrite mode
    }
    __syncwarp();
    x = warp_permute(x, lane);
    if (lane < 12) c->sponge_state[lane] = x;
    if (lane < 8) c->output_buffer[lane] = x;
    __syncwarp();
    if (lane == 0) {
        c->input_len = 0;
        c->output_len = 8;
    }
    __syncwarp();
}

// Challenger::observe_elements
__device__ __forceinline__ void observe(nlc_challenger* c, const u64* elems, size_t n, int lane) {
    for (size_t i = 0; i < n; i++) {
        if (lane == 0) {
            c->output_len = 0;
            c->input_buffer[c->input_len] = gl::canon(elems[i]);
            c->input_len++;
        }
        __syncwarp();
        if (c->input_len == 8) duplex(c, lane);
    }
}

// Challenger::get_challenge (returned to every lane)
__device__ __forceinline__ u64 get(nlc_challenger* c, int lane) {
    __syncwarp();
    if (c->input_len != 0 || c->output_len == 0) duplex(c, lane);
    u64 v = c->output_buffer[c->output_len - 1];
    __syncwarp();
    if (lane == 0) c->output_len--;
    __syncwarp();
    return v;
}

} // namespace transcript
