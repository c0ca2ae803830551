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
// (canonical); lanes 12..15 of each half idle along.  All 32 lanes must call; at most 8 warps per block.  The
// latency of ONE permutation is what this form is for (transcript, narrow Merkle layers): with a single warp on the
// scheduler every dependent instruction costs its full latency, so the round is kept short -- S-boxes run across
// lanes, the state is exchanged through shared memory (one 8-byte store, six 16-byte broadcast loads; double
// buffered, one __syncwarp per round) and lane r forms its MDS row with 24 IMAD.WIDE on coefficients it rotated
// into registers once, in four independent accumulation chains.
// Lane-varying reads of a __constant__ array are serialised per distinct address (12 replays per round here, the
// single biggest latency item of the first version): the round constants of this form live in global memory instead
// (L1-resident after the first permutation) and the next round's is fetched a round ahead.
static __device__ const uint64_t g_rc[NLC_POSEIDON_N_RC + 12] = NLC_POSEIDON_RC_INIT;
static __constant__ uint32_t c_circ2[24] = {17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20,
                                            17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20};
__device__ __forceinline__ u64 warp_permute(u64 x, int lane) {
    __shared__ ulonglong2 xch[8][2][2][6]; // [warp][half-warp][buffer][6 pairs = 12 lanes]
    const int l16 = lane & 15, l12 = l16 < 12 ? l16 : 0;
    ulonglong2(*buf)[6] = xch[(threadIdx.x >> 5) & 7][lane >> 4];
    // out[r] = sum_j x[j] * CIRC[(j - r) mod 12]  (+ 8 x[0] for r = 0: MDS_MATRIX_DIAG)
    u32 coef[12];
#pragma unroll
    for (int j = 0; j < 12; j++) coef[j] = c_circ2[j + 12 - l12];
    if (l12 == 0) coef[0] += 8;
    u64 rc = __ldg(g_rc + l12);
#pragma unroll 1
    for (int r = 0; r < 30; r++) {
        x = gl::add_nc(x, rc);
        rc = __ldg(g_rc + (r + 1) * 12 + l12); // round 30 reads the zero padding
        const bool full = r < 4 || r >= 26;
        if (full || l16 == 0) x = poseidon::sbox7(x);
        u64* mine = reinterpret_cast<u64*>(buf[r & 1]);
        if (l16 < 12) mine[l16] = x;
        __syncwarp();
        u64 a0 = 0, a1 = 0, b0 = 0, b1 = 0; // low-word and high-word sums, two chains each
#pragma unroll
        for (int q = 0; q < 6; q++) {
            const ulonglong2 v = buf[r & 1][q];
            a0 += (u64)(u32)v.x * coef[2 * q];
            b0 += (u64)(u32)(v.x >> 32) * coef[2 * q];
            a1 += (u64)(u32)v.y * coef[2 * q + 1];
            b1 += (u64)(u32)(v.y >> 32) * coef[2 * q + 1];
        }
        const u64 alo = a0 + a1, ahi = b0 + b1; // < 2^41 each
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
    size_t i = 0;
    for (; i < n; i++) {
        // Whole blocks while the input buffer is empty (an opening set is thousands of elements): the eight elements go
        // straight into the rate part of the state, which stays in registers from one permutation to the next, and the
        // next block is loaded while this one is permuted -- the element-by-element path below costs ~1.8 us of global-
        // memory bookkeeping per block on top of the 9 us permutation.  Leaves the challenger exactly as that path would:
        // state and output buffer after the last duplexing, input empty.
        if (n - i >= 8 && c->input_len == 0) {
            u64 x = lane < 12 ? c->sponge_state[lane] : 0;
            u64 nxt = lane < 8 ? elems[i + lane] : 0;
            for (; n - i >= 8; i += 8) {
                if (lane < 8) x = gl::canon(nxt);
                if (lane < 8 && n - i >= 16) nxt = elems[i + 8 + lane];
                x = warp_permute(x, lane);
            }
            if (lane < 12) c->sponge_state[lane] = x;
            if (lane < 8) c->output_buffer[lane] = x;
            __syncwarp();
            if (lane == 0) {
                c->input_len = 0;
                c->output_len = 8;
            }
            __syncwarp();
            if (i >= n) break;
        }
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
