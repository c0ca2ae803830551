/*
 * oracle/quotient.c -- the constraint-free parts of Starky's quotient_polys, CPU restatement.
 * TEST INFRASTRUCTURE ONLY.
 *
 * Follows /root/reference/vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:
 *   278-283  lagrange_first / lagrange_last = PolynomialValues::selector(degree, 0 | degree-1)
 *                                             .lde_onto_coset(quotient_degree_bits)
 *   285      ZeroPolyOnCoset::new(degree_bits, quotient_degree_bits)   (plonky2 field/zero_poly_coset.rs, source
 *            absent: evals[i] = g^n * w_{2^rate}^i - 1 over the two-adic subgroup of order 2^rate in natural
 *            order, inverses by batch inversion; eval_inverse(i) = inverses[i % 2^rate])
 *   298-302  coset = cyclic_subgroup_coset_known_order(w_size, g, size)
 *   336-340  constraints_evals *= z_h_on_coset.eval_inverse(i)
 *   350-357  transpose(quotient_values) -> per challenge values.coset_ifft(g)
 *   158-178  trim_to_len(degree * quotient_degree_factor) ("Quotient has failed ..."), chunks(degree),
 *            PolynomialBatch::from_coeffs(all_quotient_chunks, rate_bits, false, cap_height, ..)
 * The AIR evaluation itself (prover.rs:320-334) is outside the path.
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

void orc_quotient_domain(int log_n, int qdb, uint64_t* coset_x, uint64_t* lagrange_first, uint64_t* lagrange_last,
                         uint64_t* zh_inv) {
    size_t n = (size_t)1 << log_n, size = n << qdb;
    if (zh_inv) {
        gl_t g_pow_n = GL_GENERATOR;
        for (int i = 0; i < log_n; i++) g_pow_n = gl_sqr(g_pow_n); /* coset_shift().exp_power_of_2(n_log) */
        gl_t w = gl_root_of_unity(qdb), x = 1;
        for (size_t i = 0; i < ((size_t)1 << qdb); i++) {
            zh_inv[i] = gl_inv(gl_sub(gl_mul(g_pow_n, x), 1));
            x = gl_mul(x, w);
        }
    }
    if (coset_x) {
        gl_t w = gl_root_of_unity(log_n + qdb), x = GL_GENERATOR;
        for (size_t i = 0; i < size; i++) {
            coset_x[i] = x;
            x = gl_mul(x, w);
        }
    }
    for (int which = 0; which < 2; which++) {
        uint64_t* out = which ? lagrange_last : lagrange_first;
        if (!out) continue;
        uint64_t* sel = (uint64_t*)calloc(n, 8);
        sel[which ? n - 1 : 0] = 1;
        orc_lde_onto_coset(sel, log_n, qdb, out);
        free(sel);
    }
}

/* evals: size x n_ch row-major (quotient_values before transpose).  Outputs as orc_commit_from_coeffs for
 * W = n_ch * qdf columns; coeffs_out (W x n) receives the chunks.  Returns 0, or 1 when trim_to_len fails. */
int orc_quotient_commit(const uint64_t* evals, size_t n_ch, int log_n, int qdb, size_t qdf, int divide_by_zh,
                        int rate_bits, int cap_height, uint64_t* coeffs_out, uint64_t* leaves, uint64_t* digests,
                        uint64_t* cap) {
    size_t n = (size_t)1 << log_n, size = n << qdb, keep = n * qdf;
    uint64_t* zh = (uint64_t*)malloc(sizeof(uint64_t) << qdb);
    orc_quotient_domain(log_n, qdb, NULL, NULL, NULL, zh);
    uint64_t* col = (uint64_t*)malloc(size * 8);
    int bad = 0;
    for (size_t j = 0; j < n_ch; j++) {
        for (size_t i = 0; i < size; i++) {
            gl_t v = gl_canon(evals[i * n_ch + j]);
            col[i] = divide_by_zh ? gl_mul(v, zh[i & (((size_t)1 << qdb) - 1)]) : v;
        }
        orc_coset_ifft(col, log_n + qdb, GL_GENERATOR);
        for (size_t i = keep; i < size; i++) bad |= col[i] != 0;
        memcpy(coeffs_out + j * keep, col, keep * 8);
    }
    free(col);
    free(zh);
    if (bad) return 1;
    orc_commit_from_coeffs(coeffs_out, n_ch * qdf, log_n, rate_bits, cap_height, leaves, digests, cap);
    return 0;
}
