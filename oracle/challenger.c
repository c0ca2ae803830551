/*
 * oracle/challenger.c -- Poseidon duplex Fiat-Shamir transcript, CPU restatement.
 * TEST INFRASTRUCTURE ONLY.
 *
 * Restates plonky2 0.1.4 @ d2598bd plonky2/src/iop/challenger.rs (Challenger::{new, observe_element,
 * observe_elements, get_challenge, get_n_challenges, get_extension_challenge, duplexing}); source
 * absent from /root/reference.  Order of use in tree:
 * vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:67,94-100,181-183,193.
 */
#include "oracle.h"
#include <string.h>

void orc_challenger_init(orc_challenger* c) { memset(c, 0, sizeof *c); }

static void duplexing(orc_challenger* c) {
    for (int i = 0; i < c->n_input; i++) c->state[i] = c->input[i]; /* overwrite mode */
    c->n_input = 0;
    orc_poseidon_permute(c->state);
    memcpy(c->output, c->state, 8 * sizeof(uint64_t));
    c->n_output = 8;
}

void orc_challenger_observe(orc_challenger* c, const uint64_t* elems, size_t n) {
    for (size_t i = 0; i < n; i++) {
        c->n_output = 0; /* any buffered outputs are now invalid */
        c->input[c->n_input++] = gl_canon(elems[i]);
        if (c->n_input == 8) duplexing(c);
    }
}

uint64_t orc_challenger_get(orc_challenger* c) {
    if (c->n_input > 0 || c->n_output == 0) duplexing(c);
    return c->output[--c->n_output]; /* Vec::pop: from the end */
}

void orc_challenger_get_n(orc_challenger* c, size_t n, uint64_t* out) {
    for (size_t i = 0; i < n; i++) out[i] = orc_challenger_get(c);
}

void orc_challenger_get_ext(orc_challenger* c, uint64_t* out2) { orc_challenger_get_n(c, 2, out2); }
