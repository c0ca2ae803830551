/*
 * oracle/gl.h -- Goldilocks field (p = 2^64 - 2^32 + 1) and its quadratic extension, CPU restatement.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may build, link or
 * call it.  The product (near-light-client_b200/csrc) never includes this header.
 *
 * Restates plonky2_field 0.1.1 @ d2598bd (pinned by /root/reference/Cargo.lock:4912-4924; the crate
 * source is NOT under /root/reference -- SURVEY.md section 0 fact 1), files
 * field/src/goldilocks_field.rs (reduce128, add/sub/mul, MULTIPLICATIVE_GROUP_GENERATOR = 7,
 * POWER_OF_TWO_GENERATOR = 1753635133440165772, TWO_ADICITY = 32) and
 * field/src/extension/quadratic.rs + goldilocks_extensions.rs (F_p[X]/(X^2 - 7)).
 * Call sites that fix the semantics in tree: vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:185
 * (F::primitive_root_of_unity), prover.rs:356 (F::coset_shift()).
 *
 * All values crossing a function boundary here are canonical (< p).
 */
#ifndef ORACLE_GL_H
#define ORACLE_GL_H
#include <stdint.h>
#include <stddef.h>

typedef uint64_t gl_t;
#define GL_P 0xFFFFFFFF00000001ULL
#define GL_EPS 0xFFFFFFFFULL
#define GL_GENERATOR 7ULL                        /* MULTIPLICATIVE_GROUP_GENERATOR == coset_shift() */
#define GL_POW2_GENERATOR 1753635133440165772ULL /* 7^((p-1)/2^32), order exactly 2^32 */
#define GL_TWO_ADICITY 32

/* All selects are written as masks: the conditions are data dependent (about 50/50), and branches on them
 * mispredict their way to a 10x slower permutation. */
static inline gl_t gl_canon(gl_t x) { return x - (GL_P & -(uint64_t)(x >= GL_P)); }

static inline gl_t gl_add(gl_t a, gl_t b) {
    gl_t s = a + b;
    /* a,b < p so a+b < 2p: one subtraction is enough */
    return s - (GL_P & -(uint64_t)((s < a) | (s >= GL_P)));
}
static inline gl_t gl_sub(gl_t a, gl_t b) { return a - b + (GL_P & -(uint64_t)(a < b)); }
static inline gl_t gl_neg(gl_t a) { return a ? GL_P - a : 0; }

/* x = lo + 2^64 hi_lo + 2^96 hi_hi  ==  lo + eps*hi_lo - hi_hi  (2^64 == eps, 2^96 == -1 mod p) */
static inline gl_t gl_reduce128(__uint128_t x) {
    uint64_t lo = (uint64_t)x, hi = (uint64_t)(x >> 64);
    uint64_t hi_hi = hi >> 32, hi_lo = hi & GL_EPS;
    uint64_t t0 = lo - hi_hi;
    t0 -= GL_EPS & -(uint64_t)(lo < hi_hi);
    uint64_t t1 = hi_lo * GL_EPS;
    uint64_t r = t0 + t1;
    r += GL_EPS & -(uint64_t)(r < t0);
    return gl_canon(r);
}
static inline gl_t gl_mul(gl_t a, gl_t b) { return gl_reduce128((__uint128_t)a * b); }
static inline gl_t gl_sqr(gl_t a) { return gl_mul(a, a); }

static inline gl_t gl_pow(gl_t b, uint64_t e) {
    gl_t r = 1;
    while (e) {
        if (e & 1) r = gl_mul(r, b);
        b = gl_sqr(b);
        e >>= 1;
    }
    return r;
}
static inline gl_t gl_inv(gl_t a) { return gl_pow(a, GL_P - 2); }

/* primitive_root_of_unity(n_log): POWER_OF_TWO_GENERATOR squared (32 - n_log) times */
static inline gl_t gl_root_of_unity(int n_log) {
    gl_t r = GL_POW2_GENERATOR;
    for (int i = n_log; i < GL_TWO_ADICITY; i++) r = gl_sqr(r);
    return r;
}

/* ---- quadratic extension F_p[X]/(X^2 - 7); element = {a0, a1}, flattened [a0, a1] ---- */
typedef struct { gl_t c[2]; } gl2_t;
#define GL2_W 7ULL

static inline gl2_t gl2_make(gl_t a0, gl_t a1) { gl2_t r = {{a0, a1}}; return r; }
static inline gl2_t gl2_from_base(gl_t a) { return gl2_make(a, 0); }
static inline gl2_t gl2_add(gl2_t a, gl2_t b) { return gl2_make(gl_add(a.c[0], b.c[0]), gl_add(a.c[1], b.c[1])); }
static inline gl2_t gl2_sub(gl2_t a, gl2_t b) { return gl2_make(gl_sub(a.c[0], b.c[0]), gl_sub(a.c[1], b.c[1])); }
static inline gl2_t gl2_mul(gl2_t a, gl2_t b) {
    gl_t c0 = gl_add(gl_mul(a.c[0], b.c[0]), gl_mul(GL2_W, gl_mul(a.c[1], b.c[1])));
    gl_t c1 = gl_add(gl_mul(a.c[0], b.c[1]), gl_mul(a.c[1], b.c[0]));
    return gl2_make(c0, c1);
}
static inline gl2_t gl2_scalar_mul(gl2_t a, gl_t s) { return gl2_make(gl_mul(a.c[0], s), gl_mul(a.c[1], s)); }
static inline int gl2_eq(gl2_t a, gl2_t b) { return a.c[0] == b.c[0] && a.c[1] == b.c[1]; }
static inline gl2_t gl2_inv(gl2_t a) {
    /* 1/(a0 + a1 X) = (a0 - a1 X) / (a0^2 - 7 a1^2) */
    gl_t norm = gl_sub(gl_sqr(a.c[0]), gl_mul(GL2_W, gl_sqr(a.c[1])));
    gl_t ni = gl_inv(norm);
    return gl2_make(gl_mul(a.c[0], ni), gl_mul(gl_neg(a.c[1]), ni));
}
static inline gl2_t gl2_pow(gl2_t b, uint64_t e) {
    gl2_t r = gl2_from_base(1);
    while (e) {
        if (e & 1) r = gl2_mul(r, b);
        b = gl2_mul(b, b);
        e >>= 1;
    }
    return r;
}

static inline size_t bitrev(size_t x, int bits) {
    size_t r = 0;
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1) << (bits - 1 - i);
    return r;
}
#endif
