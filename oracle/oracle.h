/*
 * oracle/oracle.h -- C API of the CPU oracle for the commitment hot path
 * (Goldilocks coset-LDE -> Poseidon Merkle commit -> FRI fold/commit/query).
 *
 * TEST INFRASTRUCTURE ONLY.  Built by oracle/Makefile into oracle/liboracle.so; loaded by
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs -- never by
 * the product path.
 *
 * PARITY PIN: the Poseidon permutation is pinned by upstream plonky2's own known-answer vectors
 * (orc_poseidon_selftest).  Above the permutation the reference repository holds NO golden vector
 * for this path and plonky2 itself cannot be built here (no Rust
 * toolchain, crate source absent): sponge / FFT / Merkle / commit / FRI levels are "parity
 * unpinned" by reference data and are instead pinned by algebraic identities and by the restated
 * FRI verifier accepting the oracle's own proofs (orc_fri_verify).
 */
#ifndef ORACLE_H
#define ORACLE_H
#include "gl.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- poseidon.c ---- */
void orc_poseidon_init(void);
const uint64_t* orc_poseidon_round_constants(void); /* 360 */
int orc_poseidon_selftest(void);
void orc_poseidon_permute(uint64_t* state12);       /* fast partial rounds (what plonky2 runs) */
void orc_poseidon_permute_naive(uint64_t* state12); /* the definition: dense MDS every round */
void orc_hash_no_pad(const uint64_t* in, size_t len, uint64_t* out4);
void orc_two_to_one(const uint64_t* l4, const uint64_t* r4, uint64_t* out4);
void orc_hash_or_noop(const uint64_t* in, size_t len, uint64_t* out4);

/* ---- fft.c : natural order in, natural order out, in place ---- */
void orc_fft(uint64_t* a, int log_n);
void orc_ifft(uint64_t* a, int log_n);
void orc_coset_fft(uint64_t* a, int log_n, uint64_t shift);
void orc_coset_ifft(uint64_t* a, int log_n, uint64_t shift);
/* values (n) -> evaluations on shift*<w_N>, N = n << rate_bits (ifft, zero-pad, coset_fft) */
void orc_lde_onto_coset(const uint64_t* values, int log_n, int rate_bits, uint64_t* out);
void orc_reverse_index_bits(uint64_t* a, int log_n, size_t elem_words);

/* ---- merkle.c ---- */
/* leaves: N x leaf_len row-major.  digests: 2*(N - 2^cap) x 4 in plonky2's interleaved layout.
 * cap: 2^cap x 4. */
void orc_merkle_new(const uint64_t* leaves, size_t n_leaves, size_t leaf_len, int cap_height,
                    uint64_t* digests, uint64_t* cap);
/* siblings: (log2 N - cap_height) x 4, bottom-up */
void orc_merkle_prove(const uint64_t* digests, size_t n_leaves, int cap_height, size_t leaf_index,
                      uint64_t* siblings);
int orc_merkle_verify(const uint64_t* leaf, size_t leaf_len, size_t leaf_index, const uint64_t* cap,
                      int cap_height, const uint64_t* siblings, int n_siblings);

/* ---- commit.c : PolynomialBatch::from_values / from_coeffs ---- */
/* values/coeffs: W x n (column j contiguous).  leaves: N x W, row k = LDE row bitrev_N(k). */
void orc_commit_from_values(const uint64_t* values, size_t W, int log_n, int rate_bits, int cap_height,
                            uint64_t* coeffs, uint64_t* leaves, uint64_t* digests, uint64_t* cap);
void orc_commit_from_coeffs(const uint64_t* coeffs, size_t W, int log_n, int rate_bits, int cap_height,
                            uint64_t* leaves, uint64_t* digests, uint64_t* cap);

/* ---- quotient.c : quotient_polys without the AIR evaluation (starkyx prover.rs:253-358, 158-178) ---- */
/* tables over the quotient domain (size = n << qdb); any output may be NULL; zh_inv has 2^qdb entries */
void orc_quotient_domain(int log_n, int qdb, uint64_t* coset_x, uint64_t* lagrange_first, uint64_t* lagrange_last,
                         uint64_t* zh_inv);
/* evals: size x n_ch row-major.  coeffs_out: (n_ch*qdf) x n chunks; leaves/digests/cap as orc_commit_from_coeffs.
 * Returns 1 when trim_to_len(n * qdf) fails (a coefficient beyond the degree bound is non-zero). */
int orc_quotient_commit(const uint64_t* evals, size_t n_ch, int log_n, int qdb, size_t qdf, int divide_by_zh,
                        int rate_bits, int cap_height, uint64_t* coeffs_out, uint64_t* leaves, uint64_t* digests,
                        uint64_t* cap);

/* ---- challenger.c : iop/challenger.rs ---- */
typedef struct {
    uint64_t state[12];
    uint64_t input[8];
    int n_input;
    uint64_t output[8];
    int n_output;
} orc_challenger;
void orc_challenger_init(orc_challenger* c);
void orc_challenger_observe(orc_challenger* c, const uint64_t* elems, size_t n);
uint64_t orc_challenger_get(orc_challenger* c);
void orc_challenger_get_n(orc_challenger* c, size_t n, uint64_t* out);
void orc_challenger_get_ext(orc_challenger* c, uint64_t* out2);

/* ---- fri.c ---- */
typedef struct {
    int degree_bits;      /* log2 n of the committed polynomials */
    int rate_bits;
    int cap_height;
    int pow_bits;
    int num_queries;
    int n_rounds;         /* len(reduction_arity_bits) */
    int arity_bits[16];
} orc_fri_params;
/* FriConfig::fri_params with FriReductionStrategy::ConstantArityBits(arity_bits, final_poly_bits) */
void orc_fri_params_constant_arity(orc_fri_params* p, int degree_bits, int rate_bits, int cap_height,
                                   int pow_bits, int num_queries, int arity_bits, int final_poly_bits);

typedef struct { /* one committed oracle (a PolynomialBatch), host arrays */
    size_t W;
    const uint64_t* coeffs;  /* W x n natural order */
    const uint64_t* leaves;  /* N x W */
    const uint64_t* digests; /* interleaved layout */
    const uint64_t* cap;
} orc_batch;

typedef struct { /* FriInstanceInfo: batches of (point, [(oracle, poly)]) */
    int n_batches;
    const uint64_t* points;        /* n_batches x 2 (extension elements) */
    const uint32_t* batch_offsets; /* n_batches + 1 into poly_oracle / poly_index */
    const uint32_t* poly_oracle;
    const uint32_t* poly_index;
} orc_fri_instance;

/* Flat proof layout (u64 words), identical in the product's C ABI (include/nlc_commit.h):
 *   caps        : n_rounds x 2^cap_height x 4
 *   final_poly  : (n >> sum(arity_bits)) x 2            (n = 2^degree_bits)
 *   pow_witness : 1
 *   queries     : num_queries x { per oracle o: leaf (W_o) then siblings ((logN - cap) x 4);
 *                                 per round i: evals (2 << arity_i) then siblings ((logN_i - arity_i - cap) x 4) }
 */
size_t orc_fri_proof_words(const orc_fri_params* p, const orc_batch* oracles, int n_oracles);
/* PolynomialBatch::prove_openings.  pow_mode: 0 = smallest valid nonce (deterministic; what the GPU
 * path returns), otherwise use *forced_pow as the witness (asserted valid). Returns 0 on success. */
int orc_prove_openings(const orc_fri_instance* inst, const orc_batch* oracles, int n_oracles,
                       orc_challenger* ch, const orc_fri_params* p, uint64_t* proof_out);
/* verify_fri_proof restated (fri/verifier.rs).  openings: per batch, the claimed values of each
 * polynomial at the batch point (extension elements, concatenated in instance order).
 * oracle_caps: per oracle its cap; oracle_W: per oracle its width.  `ch` must be in the state
 * right after observe_openings.  Returns 0 if the proof verifies, else a nonzero step code. */
int orc_fri_verify(const orc_fri_instance* inst, const uint64_t* openings, const uint64_t* const* oracle_caps,
                   const size_t* oracle_W, int n_oracles, orc_challenger* ch, const orc_fri_params* p,
                   const uint64_t* proof);

/* helpers used by tests and by the STARK-shaped driver */
void orc_eval_poly_ext(const uint64_t* coeffs, size_t n, const uint64_t* z2, uint64_t* out2);
void orc_set_threads(int n);
int orc_get_threads(void);

/* ---- vec.c: element-wise field arithmetic over arrays (the packed-field ops of an AIR parser) ---- */
void orc_vec_add(const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n);
void orc_vec_sub(const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n);
void orc_vec_mul(const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n);
void orc_vec_neg(const uint64_t* a, uint64_t* out, size_t n);
void orc_vec_mul_scalar_add(const uint64_t* a, uint64_t s, const uint64_t* c, uint64_t* out, size_t n);
void orc_vec_inv(const uint64_t* a, uint64_t* out, size_t n);

#ifdef __cplusplus
}
#endif
#endif
