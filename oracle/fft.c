/*
 * oracle/fft.c -- radix-2 NTT over Goldilocks, CPU restatement.  TEST INFRASTRUCTURE ONLY.
 *
 * Restates plonky2_field 0.1.1 @ d2598bd (source absent from /root/reference):
 *   field/src/fft.rs              fft_dispatch / fft_classic: bit-reverse the input, then DIT stages
 *                                 driven by a root table; ifft_with_options = forward transform,
 *                                 then c[0]=b[0]/n, c[n/2]=b[n/2]/n, c[i]=b[n-i]/n
 *   field/src/polynomial/mod.rs   PolynomialCoeffs::{lde, coset_fft}, PolynomialValues::{ifft,
 *                                 coset_ifft, lde_onto_coset}
 * In-tree call sites: vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:280-284 (lde_onto_coset),
 * prover.rs:356 (coset_ifft(F::coset_shift())).
 * Upstream's zero_factor shortcut (skipping the first rate_bits stages of a zero-padded input) is a
 * speed-up with identical output; not restated.
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

static gl_t* root_tables[GL_TWO_ADICITY + 1]; /* root_tables[k][j] = w_{2^k}^j, j < 2^(k-1) */

static const gl_t* roots_for(int log_n) {
    gl_t* t;
#pragma omp critical(orc_roots)
    {
        t = root_tables[log_n];
        if (!t && log_n >= 1) {
            size_t half = (size_t)1 << (log_n - 1);
            t = (gl_t*)malloc(half * sizeof(gl_t));
            gl_t w = gl_root_of_unity(log_n), x = 1;
            for (size_t j = 0; j < half; j++) { t[j] = x; x = gl_mul(x, w); }
            root_tables[log_n] = t;
        }
    }
    return t;
}

void orc_reverse_index_bits(uint64_t* a, int log_n, size_t elem_words) {
    size_t n = (size_t)1 << log_n;
    uint64_t* tmp = (uint64_t*)malloc(elem_words * sizeof(uint64_t));
    for (size_t i = 0; i < n; i++) {
        size_t j = bitrev(i, log_n);
        if (i < j) {
            memcpy(tmp, a + i * elem_words, elem_words * 8);
            memcpy(a + i * elem_words, a + j * elem_words, elem_words * 8);
            memcpy(a + j * elem_words, tmp, elem_words * 8);
        }
    }
    free(tmp);
}

void orc_fft(uint64_t* a, int log_n) {
    size_t n = (size_t)1 << log_n;
    if (log_n == 0) { a[0] = gl_canon(a[0]); return; }
    const gl_t* roots = roots_for(log_n);
    for (size_t i = 0; i < n; i++) a[i] = gl_canon(a[i]);
    orc_reverse_index_bits(a, log_n, 1);
    for (int s = 1; s <= log_n; s++) {
        size_t m = (size_t)1 << s, half = m >> 1, step = n >> s;
        for (size_t k = 0; k < n; k += m) {
            for (size_t j = 0; j < half; j++) {
                gl_t u = a[k + j], v = gl_mul(a[k + j + half], roots[j * step]);
                a[k + j] = gl_add(u, v);
                a[k + j + half] = gl_sub(u, v);
            }
        }
    }
}

void orc_ifft(uint64_t* a, int log_n) {
    size_t n = (size_t)1 << log_n;
    orc_fft(a, log_n);
    gl_t n_inv = gl_inv((gl_t)(n % GL_P));
    a[0] = gl_mul(a[0], n_inv);
    if (n > 1) a[n / 2] = gl_mul(a[n / 2], n_inv);
    for (size_t i = 1; i < n / 2; i++) {
        gl_t x = gl_mul(a[i], n_inv), y = gl_mul(a[n - i], n_inv);
        a[i] = y;
        a[n - i] = x;
    }
}

void orc_coset_fft(uint64_t* a, int log_n, uint64_t shift) {
    size_t n = (size_t)1 << log_n;
    gl_t x = 1;
    shift = gl_canon(shift);
    for (size_t j = 0; j < n; j++) { a[j] = gl_mul(gl_canon(a[j]), x); x = gl_mul(x, shift); }
    orc_fft(a, log_n);
}

void orc_coset_ifft(uint64_t* a, int log_n, uint64_t shift) {
    size_t n = (size_t)1 << log_n;
    orc_ifft(a, log_n);
    gl_t si = gl_inv(gl_canon(shift)), x = 1;
    for (size_t j = 0; j < n; j++) { a[j] = gl_mul(a[j], x); x = gl_mul(x, si); }
}

void orc_lde_onto_coset(const uint64_t* values, int log_n, int rate_bits, uint64_t* out) {
    size_t n = (size_t)1 << log_n, N = n << rate_bits;
    memcpy(out, values, n * 8);
    orc_ifft(out, log_n);
    memset(out + n, 0, (N - n) * 8);
    orc_coset_fft(out, log_n + rate_bits, GL_GENERATOR);
}

/* PolynomialCoeffs<F::Extension>::eval for base-field coefficients lifted to the extension
 * (proof.rs:307-310: p.to_extension().eval(z)) -- Horner from the top coefficient. */
void orc_eval_poly_ext(const uint64_t* coeffs, size_t n, const uint64_t* z2, uint64_t* out2) {
    gl2_t z = gl2_make(gl_canon(z2[0]), gl_canon(z2[1])), acc = gl2_from_base(0);
    for (size_t i = n; i-- > 0;) acc = gl2_add(gl2_mul(acc, z), gl2_from_base(gl_canon(coeffs[i])));
    out2[0] = acc.c[0];
    out2[1] = acc.c[1];
}
