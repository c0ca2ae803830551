/*
 * oracle/vec.c -- element-wise Goldilocks arithmetic over arrays.  TEST INFRASTRUCTURE ONLY.
 *
 * The arithmetic a packed-field AIR parser performs (vendor/starkyx/starkyx/src/plonky2/parser/mod.rs:88-127:
 * StarkParser::{add, sub, neg, mul, mul_const, add_const} on P = PackedField over GoldilocksField), applied to
 * ALL points of the quotient domain at once: one array element per point.  oracle/air.py builds the direct
 * (non-bytecode) evaluation of starkyx's quotient loop (prover.rs:306-352) from these.
 */
#include "oracle.h"

void orc_vec_add(const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n) {
#pragma omp parallel for num_threads(orc_get_threads()) schedule(static)
    for (size_t i = 0; i < n; i++) out[i] = gl_add(gl_canon(a[i]), gl_canon(b[i]));
}
void orc_vec_sub(const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n) {
#pragma omp parallel for num_threads(orc_get_threads()) schedule(static)
    for (size_t i = 0; i < n; i++) out[i] = gl_sub(gl_canon(a[i]), gl_canon(b[i]));
}
void orc_vec_mul(const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n) {
#pragma omp parallel for num_threads(orc_get_threads()) schedule(static)
    for (size_t i = 0; i < n; i++) out[i] = gl_mul(gl_canon(a[i]), gl_canon(b[i]));
}
void orc_vec_neg(const uint64_t* a, uint64_t* out, size_t n) {
#pragma omp parallel for num_threads(orc_get_threads()) schedule(static)
    for (size_t i = 0; i < n; i++) out[i] = gl_neg(gl_canon(a[i]));
}
/* out = a * s + c   (the ConstraintConsumer update acc = acc * alpha + constraint, consumer.rs:60-66) */
void orc_vec_mul_scalar_add(const uint64_t* a, uint64_t s, const uint64_t* c, uint64_t* out, size_t n) {
    s = gl_canon(s);
#pragma omp parallel for num_threads(orc_get_threads()) schedule(static)
    for (size_t i = 0; i < n; i++) out[i] = gl_add(gl_mul(gl_canon(a[i]), s), gl_canon(c[i]));
}
/* element-wise inverse (Fermat); 0 has none: plonky2's batch_multiplicative_inverse panics on it, here out = 0 */
void orc_vec_inv(const uint64_t* a, uint64_t* out, size_t n) {
#pragma omp parallel for num_threads(orc_get_threads()) schedule(static)
    for (size_t i = 0; i < n; i++) out[i] = gl_inv(gl_canon(a[i]));
}
