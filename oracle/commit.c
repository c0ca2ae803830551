/*
 * oracle/commit.c -- PolynomialBatch::from_values / from_coeffs, CPU restatement.
 * TEST INFRASTRUCTURE ONLY.
 *
 * Restates plonky2 0.1.4 @ d2598bd plonky2/src/fri/oracle.rs (from_values, from_coeffs, lde_values;
 * source absent from /root/reference) with blinding = false, as called at
 * vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:91-93 (from_values), prover.rs:171-178
 * (from_coeffs) and plonky2/stark/config.rs:92-94:
 *   "IFFT"             coeffs_j = ifft(values_j)                         (parallel over columns)
 *   "FFT + blinding"   lde_j = coset_fft(zero-pad(coeffs_j, N), shift 7) (parallel over columns)
 *   "transpose LDEs"   leaves[k] = [lde_0[bitrev_N(k)], ..., lde_{W-1}[bitrev_N(k)]]
 *   "build Merkle tree" MerkleTree::new(leaves, cap_height)
 * OpenMP mirrors rayon's decomposition so the timing is a fair multi-core CPU baseline.
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

void orc_commit_from_coeffs(const uint64_t* coeffs, size_t W, int log_n, int rate_bits, int cap_height,
                            uint64_t* leaves, uint64_t* digests, uint64_t* cap) {
    size_t n = (size_t)1 << log_n, N = n << rate_bits;
    int log_N = log_n + rate_bits;
    uint64_t* lde = (uint64_t*)malloc(W * N * sizeof(uint64_t));
    { /* build the root table once, outside the parallel region */
        uint64_t* warm = (uint64_t*)calloc(N, 8);
        orc_fft(warm, log_N);
        free(warm);
    }
#pragma omp parallel for num_threads(orc_get_threads()) schedule(dynamic, 1)
    for (size_t j = 0; j < W; j++) {
        uint64_t* col = lde + j * N;
        memcpy(col, coeffs + j * n, n * 8);
        memset(col + n, 0, (N - n) * 8);
        orc_coset_fft(col, log_N, GL_GENERATOR);
    }
    /* transpose + reverse_index_bits_in_place, blocked over rows */
#pragma omp parallel for num_threads(orc_get_threads()) schedule(static)
    for (size_t sb = 0; sb < N; sb += 8) { /* 8 consecutive LDE points = one cache line per column */
        size_t se = sb + 8 < N ? sb + 8 : N;
        uint64_t* rows[8];
        for (size_t s = sb; s < se; s++) rows[s - sb] = leaves + bitrev(s, log_N) * W;
        for (size_t j = 0; j < W; j++) {
            const uint64_t* src = lde + j * N + sb;
            for (size_t t = 0; t < se - sb; t++) rows[t][j] = src[t];
        }
    }
    free(lde);
    orc_merkle_new(leaves, N, W, cap_height, digests, cap);
}

void orc_commit_from_values(const uint64_t* values, size_t W, int log_n, int rate_bits, int cap_height,
                            uint64_t* coeffs, uint64_t* leaves, uint64_t* digests, uint64_t* cap) {
    size_t n = (size_t)1 << log_n;
    {
        uint64_t* warm = (uint64_t*)calloc(n, 8);
        orc_fft(warm, log_n);
        free(warm);
    }
#pragma omp parallel for num_threads(orc_get_threads()) schedule(dynamic, 1)
    for (size_t j = 0; j < W; j++) {
        memcpy(coeffs + j * n, values + j * n, n * 8);
        orc_ifft(coeffs + j * n, log_n);
    }
    orc_commit_from_coeffs(coeffs, W, log_n, rate_bits, cap_height, leaves, digests, cap);
}
