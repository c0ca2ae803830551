/*
 * oracle/fri.c -- PolynomialBatch::prove_openings, the FRI prover (commit phase, proof of work,
 * query rounds) and the FRI verifier, CPU restatement.  TEST INFRASTRUCTURE ONLY.
 *
 * Restates plonky2 0.1.4 @ d2598bd (source absent from /root/reference):
 *   plonky2/src/fri/oracle.rs               PolynomialBatch::prove_openings
 *   plonky2/src/util/reducing.rs            ReducingFactor::{reduce_polys_base, shift_poly, reduce}
 *   field/src/polynomial/division.rs        PolynomialCoeffs::divide_by_linear
 *   plonky2/src/fri/prover.rs               fri_proof, fri_committed_trees, fri_proof_of_work,
 *                                           fri_prover_query_rounds, fri_prover_query_round
 *   plonky2/src/fri/reduction_strategies.rs FriReductionStrategy::ConstantArityBits
 *   plonky2/src/fri/verifier.rs             verify_fri_proof, fri_combine_initial,
 *                                           fri_verifier_query_round, compute_evaluation
 *   plonky2/src/iop/challenger.rs           Challenger::fri_challenges
 * In-tree call site: vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:200-206 (prove_openings),
 * knobs at plonky2/stark/config.rs:64-70, verifier use at plonky2/stark/verifier.rs:148-171.
 *
 * Proof of work: upstream searches with rayon find_any (non-deterministic winner);
 * this oracle returns the SMALLEST valid nonce, which is what the GPU grinder returns too.
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

void orc_fri_params_constant_arity(orc_fri_params* p, int degree_bits, int rate_bits, int cap_height, int pow_bits,
                                   int num_queries, int arity_bits, int final_poly_bits) {
    memset(p, 0, sizeof *p);
    p->degree_bits = degree_bits; p->rate_bits = rate_bits; p->cap_height = cap_height;
    p->pow_bits = pow_bits; p->num_queries = num_queries;
    int d = degree_bits;
    while (d > final_poly_bits && d + rate_bits - arity_bits >= cap_height) {
        p->arity_bits[p->n_rounds++] = arity_bits;
        d -= arity_bits;
    }
}

static int total_arity(const orc_fri_params* p) {
    int t = 0;
    for (int i = 0; i < p->n_rounds; i++) t += p->arity_bits[i];
    return t;
}

size_t orc_fri_proof_words(const orc_fri_params* p, const orc_batch* oracles, int n_oracles) {
    int log_N = p->degree_bits + p->rate_bits;
    size_t w = (size_t)p->n_rounds * ((size_t)4 << p->cap_height);
    w += ((size_t)2 << (p->degree_bits - total_arity(p)));
    w += 1;
    size_t q = 0;
    for (int o = 0; o < n_oracles; o++) q += oracles[o].W + (size_t)4 * (log_N - p->cap_height);
    int lg = log_N;
    for (int i = 0; i < p->n_rounds; i++) {
        q += ((size_t)2 << p->arity_bits[i]) + (size_t)4 * (lg - p->arity_bits[i] - p->cap_height);
        lg -= p->arity_bits[i];
    }
    return w + q * p->num_queries;
}

/* in-place FFT over F_{p^2} with base-field roots: componentwise (fft.rs is generic over
 * Field with the roots taken from the base field, so the transform is F_p-linear) */
static void ext_coset_fft(gl2_t* a, int log_n, gl_t shift) {
    size_t n = (size_t)1 << log_n;
    uint64_t* c0 = (uint64_t*)malloc(n * 8);
    uint64_t* c1 = (uint64_t*)malloc(n * 8);
    for (size_t i = 0; i < n; i++) { c0[i] = a[i].c[0]; c1[i] = a[i].c[1]; }
    orc_coset_fft(c0, log_n, shift);
    orc_coset_fft(c1, log_n, shift);
    for (size_t i = 0; i < n; i++) a[i] = gl2_make(c0[i], c1[i]);
    free(c0);
    free(c1);
}

static int pow_ok(uint64_t response, int min_zeros) {
    /* to_canonical_u64().leading_zeros() >= proof_of_work_bits + (64 - order().bits()) = pow_bits */
    return min_zeros == 0 || (response >> (64 - min_zeros)) == 0;
}

typedef struct { uint64_t* leaves; uint64_t* digests; uint64_t* cap; size_t n_leaves; size_t leaf_len; } fri_tree;

int orc_prove_openings(const orc_fri_instance* inst, const orc_batch* oracles, int n_oracles, orc_challenger* ch,
                       const orc_fri_params* p, uint64_t* out) {
    const int log_n = p->degree_bits, log_N = log_n + p->rate_bits;
    const size_t n = (size_t)1 << log_n, N = (size_t)1 << log_N;

    uint64_t a2[2];
    orc_challenger_get_ext(ch, a2);
    const gl2_t alpha = gl2_make(a2[0], a2[1]);

    gl2_t* final_poly = (gl2_t*)calloc(N, sizeof(gl2_t)); /* lde(): zero padded to N already */
    gl2_t* comp = (gl2_t*)malloc(n * sizeof(gl2_t));
    for (int b = 0; b < inst->n_batches; b++) {
        const uint32_t lo = inst->batch_offsets[b], hi = inst->batch_offsets[b + 1];
        const gl2_t z = gl2_make(gl_canon(inst->points[2 * b]), gl_canon(inst->points[2 * b + 1]));
        /* reduce_polys_base: sum_k alpha^k * poly_k */
        memset(comp, 0, n * sizeof(gl2_t));
        gl2_t apow = gl2_from_base(1);
        for (uint32_t k = lo; k < hi; k++) {
            const uint64_t* poly = oracles[inst->poly_oracle[k]].coeffs + (size_t)inst->poly_index[k] * n;
            for (size_t i = 0; i < n; i++) comp[i] = gl2_add(comp[i], gl2_scalar_mul(apow, gl_canon(poly[i])));
            apow = gl2_mul(apow, alpha);
        }
        /* shift_poly: final_poly *= alpha^count, count = polys in this batch */
        const gl2_t shift = gl2_pow(alpha, hi - lo);
        for (size_t i = 0; i < n; i++) final_poly[i] = gl2_mul(final_poly[i], shift);
        /* divide_by_linear(z): q[i-1] = c[i] + z*q[i]; remainder dropped; then a zero is pushed */
        gl2_t acc = gl2_from_base(0);
        for (size_t i = n; i-- > 1;) {
            acc = gl2_add(gl2_mul(acc, z), comp[i]);
            final_poly[i - 1] = gl2_add(final_poly[i - 1], acc);
        }
    }
    free(comp);

    gl2_t* coeffs = final_poly; /* lde_final_poly: N coefficients, only the first n non-zero */
    gl2_t* values = (gl2_t*)malloc(N * sizeof(gl2_t));
    memcpy(values, coeffs, N * sizeof(gl2_t));
    ext_coset_fft(values, log_N, GL_GENERATOR);

    /* ---- fri_committed_trees ---- */
    fri_tree* trees = (fri_tree*)calloc(p->n_rounds ? p->n_rounds : 1, sizeof(fri_tree));
    uint64_t* w = out;
    gl_t shift = GL_GENERATOR;
    int lg = log_N;
    size_t cur = N;
    const size_t cap_words = (size_t)4 << p->cap_height;
    for (int r = 0; r < p->n_rounds; r++) {
        const int ab = p->arity_bits[r];
        const size_t arity = (size_t)1 << ab;
        orc_reverse_index_bits((uint64_t*)values, lg, 2);
        fri_tree* t = &trees[r];
        t->n_leaves = cur >> ab;
        t->leaf_len = 2 * arity;
        t->leaves = (uint64_t*)malloc(cur * 2 * 8);
        memcpy(t->leaves, values, cur * 2 * 8); /* chunks of `arity` extension elements, flattened */
        t->digests = (uint64_t*)malloc((2 * (t->n_leaves - ((size_t)1 << p->cap_height)) * 4 + 4) * 8);
        t->cap = w;
        orc_merkle_new(t->leaves, t->n_leaves, t->leaf_len, p->cap_height, t->digests, t->cap);
        orc_challenger_observe(ch, t->cap, cap_words);
        w += cap_words;
        uint64_t b2[2];
        orc_challenger_get_ext(ch, b2);
        const gl2_t beta = gl2_make(b2[0], b2[1]);
        /* coeffs'[i] = reduce_with_powers(chunk_i, beta) = sum_k chunk[k] beta^k */
        const size_t nxt = cur >> ab;
        for (size_t i = 0; i < nxt; i++) {
            gl2_t acc = gl2_from_base(0);
            for (size_t k = arity; k-- > 0;) acc = gl2_add(gl2_mul(acc, beta), coeffs[i * arity + k]);
            coeffs[i] = acc;
        }
        shift = gl_pow(shift, arity);
        cur = nxt;
        lg -= ab;
        memcpy(values, coeffs, cur * sizeof(gl2_t));
        ext_coset_fft(values, lg, shift);
    }
    /* truncate the (zero) high coefficients, observe the final polynomial */
    const size_t final_len = cur >> p->rate_bits;
    for (size_t i = 0; i < final_len; i++) { w[2 * i] = coeffs[i].c[0]; w[2 * i + 1] = coeffs[i].c[1]; }
    orc_challenger_observe(ch, w, 2 * final_len);
    w += 2 * final_len;

    /* ---- fri_proof_of_work: smallest nonce ---- */
    {
        uint64_t inter[12];
        memcpy(inter, ch->state, sizeof inter);
        for (int i = 0; i < ch->n_input; i++) inter[i] = ch->input[i];
        const int pos = ch->n_input;
        uint64_t nonce = 0;
        for (;; nonce++) {
            uint64_t s[12];
            memcpy(s, inter, sizeof s);
            s[pos] = nonce;
            orc_poseidon_permute(s);
            if (pow_ok(s[7], p->pow_bits)) break; /* squeeze().last() = state[RATE-1] */
        }
        orc_challenger_observe(ch, &nonce, 1);
        uint64_t resp = orc_challenger_get(ch);
        if (!pow_ok(resp, p->pow_bits)) return 10;
        *w++ = nonce;
    }

    /* ---- fri_prover_query_rounds ---- */
    uint64_t* rands = (uint64_t*)malloc(p->num_queries * 8);
    orc_challenger_get_n(ch, p->num_queries, rands);
    for (int q = 0; q < p->num_queries; q++) {
        size_t x = (size_t)(rands[q] % N);
        for (int o = 0; o < n_oracles; o++) {
            memcpy(w, oracles[o].leaves + x * oracles[o].W, oracles[o].W * 8);
            w += oracles[o].W;
            orc_merkle_prove(oracles[o].digests, N, p->cap_height, x, w);
            w += (size_t)4 * (log_N - p->cap_height);
        }
        int l = log_N;
        for (int r = 0; r < p->n_rounds; r++) {
            const int ab = p->arity_bits[r];
            const fri_tree* t = &trees[r];
            memcpy(w, t->leaves + (x >> ab) * t->leaf_len, t->leaf_len * 8);
            w += t->leaf_len;
            orc_merkle_prove(t->digests, t->n_leaves, p->cap_height, x >> ab, w);
            w += (size_t)4 * (l - ab - p->cap_height);
            x >>= ab;
            l -= ab;
        }
    }
    free(rands);
    for (int r = 0; r < p->n_rounds; r++) { free(trees[r].leaves); free(trees[r].digests); }
    free(trees);
    free(values);
    free(final_poly);
    return 0;
}

/* Lagrange interpolation through (x_i, y_i), evaluated at beta (the value compute_evaluation's
 * barycentric formula returns). */
static gl2_t interpolate_at(const gl_t* xs, const gl2_t* ys, size_t m, gl2_t beta) {
    gl2_t acc = gl2_from_base(0);
    for (size_t i = 0; i < m; i++) {
        gl2_t num = gl2_from_base(1);
        gl_t den = 1;
        for (size_t j = 0; j < m; j++) {
            if (j == i) continue;
            num = gl2_mul(num, gl2_sub(beta, gl2_from_base(xs[j])));
            den = gl_mul(den, gl_sub(xs[i], xs[j]));
        }
        acc = gl2_add(acc, gl2_mul(ys[i], gl2_scalar_mul(num, gl_inv(den))));
    }
    return acc;
}

int orc_fri_verify(const orc_fri_instance* inst, const uint64_t* openings, const uint64_t* const* oracle_caps,
                   const size_t* oracle_W, int n_oracles, orc_challenger* ch, const orc_fri_params* p,
                   const uint64_t* proof) {
    const int log_N = p->degree_bits + p->rate_bits;
    const size_t N = (size_t)1 << log_N;
    const size_t cap_words = (size_t)4 << p->cap_height;
    const size_t final_len = (size_t)1 << (p->degree_bits - total_arity(p));
    const uint64_t* caps = proof;
    const uint64_t* final_poly = caps + p->n_rounds * cap_words;
    const uint64_t* pow_w = final_poly + 2 * final_len;
    const uint64_t* queries = pow_w + 1;

    /* Challenger::fri_challenges */
    uint64_t t2[2];
    orc_challenger_get_ext(ch, t2);
    const gl2_t alpha = gl2_make(t2[0], t2[1]);
    gl2_t betas[16];
    for (int r = 0; r < p->n_rounds; r++) {
        orc_challenger_observe(ch, caps + r * cap_words, cap_words);
        orc_challenger_get_ext(ch, t2);
        betas[r] = gl2_make(t2[0], t2[1]);
    }
    orc_challenger_observe(ch, final_poly, 2 * final_len);
    orc_challenger_observe(ch, pow_w, 1);
    if (*pow_w >= GL_P) return 1;
    if (!pow_ok(orc_challenger_get(ch), p->pow_bits)) return 2;

    /* PrecomputedReducedOpenings: per batch sum_k alpha^k * opening_k */
    gl2_t* reduced = (gl2_t*)malloc(inst->n_batches * sizeof(gl2_t));
    for (int b = 0; b < inst->n_batches; b++) {
        gl2_t acc = gl2_from_base(0);
        for (uint32_t k = inst->batch_offsets[b + 1]; k-- > inst->batch_offsets[b];)
            acc = gl2_add(gl2_mul(acc, alpha), gl2_make(openings[2 * k], openings[2 * k + 1]));
        reduced[b] = acc;
    }

    int rc = 0;
    const uint64_t* w = queries;
    const gl_t omega_N = gl_root_of_unity(log_N);
    for (int q = 0; q < p->num_queries && !rc; q++) {
        size_t x_index = (size_t)(orc_challenger_get(ch) % N);
        /* fri_verify_initial_proof */
        const uint64_t** leaf = (const uint64_t**)malloc(n_oracles * sizeof(*leaf));
        for (int o = 0; o < n_oracles; o++) {
            leaf[o] = w;
            w += oracle_W[o];
            if (orc_merkle_verify(leaf[o], oracle_W[o], x_index, oracle_caps[o], p->cap_height, w,
                                  log_N - p->cap_height)) rc = 3;
            w += (size_t)4 * (log_N - p->cap_height);
        }
        gl_t subgroup_x = gl_mul(GL_GENERATOR, gl_pow(omega_N, bitrev(x_index, log_N)));
        /* fri_combine_initial */
        gl2_t sum = gl2_from_base(0);
        for (int b = 0; b < inst->n_batches; b++) {
            const uint32_t lo = inst->batch_offsets[b], hi = inst->batch_offsets[b + 1];
            gl2_t acc = gl2_from_base(0);
            for (uint32_t k = hi; k-- > lo;)
                acc = gl2_add(gl2_mul(acc, alpha), gl2_from_base(gl_canon(leaf[inst->poly_oracle[k]][inst->poly_index[k]])));
            gl2_t numer = gl2_sub(acc, reduced[b]);
            gl2_t denom = gl2_sub(gl2_from_base(subgroup_x), gl2_make(gl_canon(inst->points[2 * b]), gl_canon(inst->points[2 * b + 1])));
            sum = gl2_mul(sum, gl2_pow(alpha, hi - lo)); /* alpha.shift(sum) */
            sum = gl2_add(sum, gl2_mul(numer, gl2_inv(denom)));
        }
        free(leaf);
        gl2_t old_eval = sum;
        int l = log_N;
        for (int r = 0; r < p->n_rounds && !rc; r++) {
            const int ab = p->arity_bits[r];
            const size_t arity = (size_t)1 << ab;
            const uint64_t* evals = w;
            w += 2 * arity;
            const size_t coset_index = x_index >> ab, within = x_index & (arity - 1);
            if (!gl2_eq(gl2_make(evals[2 * within], evals[2 * within + 1]), old_eval)) rc = 4;
            /* compute_evaluation */
            gl_t g = gl_root_of_unity(ab);
            gl_t xs[64];
            gl2_t ys[64];
            size_t rev_within = bitrev(within, ab);
            gl_t coset_start = gl_mul(subgroup_x, gl_pow(g, arity - rev_within));
            gl_t gp = 1;
            for (size_t i = 0; i < arity; i++) {
                xs[i] = gl_mul(coset_start, gp);
                gp = gl_mul(gp, g);
                size_t src = bitrev(i, ab); /* evals reordered by reverse_index_bits */
                ys[i] = gl2_make(evals[2 * src], evals[2 * src + 1]);
            }
            old_eval = interpolate_at(xs, ys, arity, betas[r]);
            if (orc_merkle_verify(evals, 2 * arity, coset_index, caps + r * cap_words, p->cap_height, w,
                                  l - ab - p->cap_height)) rc = 5;
            w += (size_t)4 * (l - ab - p->cap_height);
            for (int i = 0; i < ab; i++) subgroup_x = gl_sqr(subgroup_x);
            x_index = coset_index;
            l -= ab;
        }
        if (!rc) { /* final_poly.eval(subgroup_x) == old_eval */
            gl2_t acc = gl2_from_base(0);
            for (size_t i = final_len; i-- > 0;)
                acc = gl2_add(gl2_scalar_mul(acc, subgroup_x), gl2_make(final_poly[2 * i], final_poly[2 * i + 1]));
            if (!gl2_eq(acc, old_eval)) rc = 6;
        }
    }
    free(reduced);
    return rc;
}
