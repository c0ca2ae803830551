/*
 * oracle/merkle.c -- MerkleTree::new / prove / verify_merkle_proof_to_cap, CPU restatement.
 * TEST INFRASTRUCTURE ONLY.
 *
 * Restates plonky2 0.1.4 @ d2598bd plonky2/src/hash/merkle_tree.rs (fill_digests_buf, fill_subtree,
 * MerkleTree::{new,prove}) and merkle_proofs.rs (verify_merkle_proof_to_cap); source absent from
 * /root/reference.  In-tree consumers: `commitment.merkle_tree.cap` at
 * vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:95,180,210.
 *
 * digests layout inside one cap subtree (recursive):
 *   [ left-subtree buffer | digest(left child) | digest(right child) | right-subtree buffer ]
 * Parallelism mirrors upstream: one task per cap subtree, rayon::join -> omp task inside.
 */
#include "oracle.h"
#include <string.h>

static int g_threads = 0;
void orc_set_threads(int n) { g_threads = n; }
#ifdef _OPENMP
#include <omp.h>
int orc_get_threads(void) { return g_threads > 0 ? g_threads : omp_get_max_threads(); }
#else
int orc_get_threads(void) { return 1; }
#endif

/* digests_buf: (2*(n_leaves-1)) hashes.  Returns the subtree root in out4. */
static void fill_subtree(uint64_t* digests_buf, size_t n_digests, const uint64_t* leaves, size_t n_leaves,
                         size_t leaf_len, uint64_t* out4) {
    if (n_digests == 0) {
        orc_hash_or_noop(leaves, leaf_len, out4);
        return;
    }
    size_t half = n_digests / 2; /* left half = [left buf (half-1) | left digest], right = [right digest | right buf] */
    uint64_t* left_buf = digests_buf;
    uint64_t* left_digest = digests_buf + (half - 1) * 4;
    uint64_t* right_digest = digests_buf + half * 4;
    uint64_t* right_buf = digests_buf + (half + 1) * 4;
    size_t hl = n_leaves / 2;
    if (n_leaves >= 1024) {
#pragma omp task default(shared)
        fill_subtree(left_buf, half - 1, leaves, hl, leaf_len, left_digest);
#pragma omp task default(shared)
        fill_subtree(right_buf, half - 1, leaves + hl * leaf_len, hl, leaf_len, right_digest);
#pragma omp taskwait
    } else {
        fill_subtree(left_buf, half - 1, leaves, hl, leaf_len, left_digest);
        fill_subtree(right_buf, half - 1, leaves + hl * leaf_len, hl, leaf_len, right_digest);
    }
    orc_two_to_one(left_digest, right_digest, out4);
}

void orc_merkle_new(const uint64_t* leaves, size_t n_leaves, size_t leaf_len, int cap_height, uint64_t* digests,
                    uint64_t* cap) {
    orc_poseidon_init();
    size_t n_cap = (size_t)1 << cap_height;
    size_t n_digests = 2 * (n_leaves - n_cap);
    if (n_digests == 0) { /* tree is all cap */
#pragma omp parallel for num_threads(orc_get_threads()) schedule(static)
        for (size_t i = 0; i < n_leaves; i++) orc_hash_or_noop(leaves + i * leaf_len, leaf_len, cap + i * 4);
        return;
    }
    size_t sub_digests = n_digests >> cap_height, sub_leaves = n_leaves >> cap_height;
#pragma omp parallel num_threads(orc_get_threads())
#pragma omp single
    for (size_t s = 0; s < n_cap; s++) {
#pragma omp task default(shared) firstprivate(s)
        fill_subtree(digests + s * sub_digests * 4, sub_digests, leaves + s * sub_leaves * leaf_len, sub_leaves,
                     leaf_len, cap + s * 4);
    }
}

void orc_merkle_prove(const uint64_t* digests, size_t n_leaves, int cap_height, size_t leaf_index,
                      uint64_t* siblings) {
    int log_n = 0;
    while (((size_t)1 << log_n) < n_leaves) log_n++;
    int num_layers = log_n - cap_height;
    size_t n_digests = 2 * (n_leaves - ((size_t)1 << cap_height));
    size_t tree_len = n_digests >> cap_height;
    const uint64_t* tree = digests + tree_len * (leaf_index >> num_layers) * 4;
    size_t pair_index = leaf_index & (((size_t)1 << num_layers) - 1);
    for (int i = 0; i < num_layers; i++) {
        size_t parity = pair_index & 1;
        pair_index >>= 1;
        size_t siblings_index = (pair_index << (i + 1)) + ((size_t)1 << i) - 1;
        size_t sibling_index = 2 * siblings_index + (1 - parity);
        memcpy(siblings + (size_t)i * 4, tree + sibling_index * 4, 32);
    }
}

int orc_merkle_verify(const uint64_t* leaf, size_t leaf_len, size_t leaf_index, const uint64_t* cap, int cap_height,
                      const uint64_t* siblings, int n_siblings) {
    (void)cap_height;
    uint64_t cur[4], nxt[4];
    orc_hash_or_noop(leaf, leaf_len, cur);
    size_t index = leaf_index;
    for (int i = 0; i < n_siblings; i++) {
        const uint64_t* sib = siblings + (size_t)i * 4;
        if (index & 1) orc_two_to_one(sib, cur, nxt); else orc_two_to_one(cur, sib, nxt);
        memcpy(cur, nxt, 32);
        index >>= 1;
    }
    return memcmp(cur, cap + index * 4, 32) == 0 ? 0 : 1;
}
