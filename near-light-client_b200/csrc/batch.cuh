// batch.cuh -- device-resident images of plonky2's PolynomialBatch and MerkleTree (internal header).
#pragma once
#include "common.cuh"
#include "gl.cuh"

struct nlc_batch {
    nlc_ctx* ctx = nullptr;
    size_t W = 0;
    int log_n = 0, rate_bits = 0, cap_height = 0;
    u64* coeffs = nullptr;  // W x n, column-major
    u64* lde = nullptr;     // W x N, column-major: lde[j*N + k] = column j of leaf k (leaf k = LDE point bitrev_N(k))
    u64* digests = nullptr; // plonky2 interleaved layout
    u64* cap = nullptr;     // 2^cap x 4
    size_t n_digest_words = 0;
    // Coset shard: this batch holds leaves [shard_idx, shard_idx+1) * (N >> shard_log) only,
    // i.e. 2^(rate_bits - shard_log) whole cosets and 2^(cap_height - shard_log) whole cap subtrees.
    int shard_log = 0;
    size_t shard_idx = 0;
    // Leaf-row shard: same leaf range as a coset shard, but the rows arrived over
    // the all-to-all: the LDE lives in the caller's receive buffer (not owned) and the coefficients stay with the
    // ranks that own the columns (coeffs == nullptr here).
    bool owns_lde = true;
    // Coefficients that live in a caller-owned buffer (the all-gathered coefficient matrix of a coset-sharded commit)
    bool owns_coeffs = true;
    size_t n_local() const { return ((size_t)1 << (log_n + rate_bits)) >> shard_log; }
    int cap_local() const { return cap_height - shard_log; }
};

struct nlc_tree {
    nlc_ctx* ctx = nullptr;
    size_t n_leaves = 0, leaf_len = 0;
    int cap_height = 0;
    u64* leaves = nullptr;
    u64* digests = nullptr;
    u64* cap = nullptr;
    bool owns_leaves = true;
};

// ---- commit.cu ----
nlc_batch* batch_alloc(nlc_ctx* ctx, size_t W, int log_n, int rate_bits, int cap_height, int shard_log = 0,
                       size_t shard_idx = 0);
void batch_release(nlc_batch* b);
void batch_ifft_values(nlc_batch* b, const u64* values_dev);
void batch_lde_and_tree(nlc_batch* b);
void batch_ifft_cols(nlc_batch* b, const u64* values_dev, size_t c0, size_t nc);
void batch_lde_cols(nlc_batch* b, size_t c0, size_t nc, const NttScatter* sc = nullptr);
void batch_tree(nlc_batch* b);
void transpose_trace_cols(nlc_ctx* ctx, const u64* rows_dev, size_t width, size_t col0, size_t n_cols, size_t n_rows,
                          u64* out_dev, size_t out_stride);
void batch_gather_lde_rows(nlc_batch* b, const u64* idx_dev, u64 i_start, size_t m, size_t step, u64* out_dev);
// true when the two device arrays differ (one small kernel + a 4-byte D2H; synchronises the stream)
bool device_words_differ(nlc_ctx* ctx, const u64* a_dev, const u64* b_dev, size_t words);
void batch_eval_ext(nlc_batch* b, const u64 z[2], u64* out_dev);
// rows [row0, row0+n_rows) of the leaf matrix, row-major, into out_dev (n_rows x W): plonky2's merkle_tree.leaves
void batch_leaf_rows(nlc_batch* b, size_t row0, size_t n_rows, u64* out_dev, cudaStream_t on = nullptr); // on: default = the context's stream

// transforms of a column shard (no tree), and the tree over a leaf-row shard
void lde_columns_dev(nlc_ctx* ctx, const u64* src_dev, size_t n_cols, int log_n, int rate_bits, bool is_values,
                     u64* coeffs_out_dev, u64* lde_out_dev, const NttScatter* sc = nullptr, size_t col0 = 0);
nlc_batch* commit_from_lde_rows_dev(nlc_ctx* ctx, u64* lde_rows_dev, size_t W, int log_n, int rate_bits, int cap_height,
                                    int shard_log, size_t shard_idx);
// iFFT of the columns one rank owns, every result stored into all n_peers coefficient matrices
void ifft_columns_bcast_dev(nlc_ctx* ctx, const u64* values_dev, size_t n_cols, int log_n, size_t col0, size_t n_peers,
                            u64* const* peer_coeffs);
// ... and this rank's coset shard of from_coeffs over the gathered matrix (read in place, not copied)
nlc_batch* commit_from_coeffs_shard_dev(nlc_ctx* ctx, u64* coeffs_dev, size_t W, int log_n, int rate_bits, int cap_height,
                                        int shard_log, size_t shard_idx);
// SURVEY.md 8f.1 (prover.rs:253-358 without the AIR evaluation)
std::vector<u64> zero_poly_on_coset_inverses(int log_n, int qdb);
void coset_points(nlc_ctx* ctx, int log_size, u64* out_dev);
nlc_batch* quotient_commit_dev(nlc_ctx* ctx, const u64* evals_dev, size_t n_ch, int log_n, int qdb, size_t qdf,
                               bool divide_by_zh, int rate_bits, int cap_height);

// ---- air.cu ----
void quotient_tables_dev(nlc_ctx* ctx, int log_n, int qdb, const u64** x, const u64** lf, const u64** ll);
void air_eval_dev(nlc_ctx* ctx, const nlc_air_program* prog, nlc_batch* const* batches, int n_batches, const u64* publics,
                  const u64* globals, const u64* challenges, const u64* alphas, size_t n_alphas, int qdb, u64* out_dev);

// ---- plonk.cu ----
nlc_batch* plonk_zs_partial_products_commit_dev(nlc_ctx* ctx, const nlc_plonk_circuit* c, nlc_batch* cs, nlc_batch* wires,
                                                const u64* betas, const u64* gammas, int cap_height);
void plonk_vanishing_dev(nlc_ctx* ctx, const nlc_plonk_circuit* c, const nlc_air_program* gates, nlc_batch* cs, nlc_batch* wires,
                         nlc_batch* zs_pp, const u64* public_inputs_hash, const u64* betas, const u64* gammas,
                         const u64* alphas, u64* out_dev);

// ---- fri.cu ----
void challenger_observe_dev(nlc_ctx* ctx, nlc_challenger* ch_dev, const u64* elems_dev, size_t n);
void challenger_get_dev(nlc_ctx* ctx, nlc_challenger* ch_dev, size_t n, u64* out_dev);
void fri_pow_dev(nlc_ctx* ctx, const nlc_challenger* ch_dev, int pow_bits, u64* witness_dev);
struct nlc_fri_state;
size_t fri_proof_words_impl(const nlc_fri_params* p, nlc_batch* const* oracles, int n_oracles, size_t* final_off,
                            size_t* pow_off, size_t* query_off, size_t* query_words);
nlc_fri_state* fri_state_new(nlc_ctx* ctx, const nlc_fri_instance* inst, nlc_batch* const* oracles, int n_oracles,
                             const nlc_fri_params* p);
void fri_state_release(nlc_fri_state* s);
void fri_begin(nlc_fri_state* s, const nlc_fri_instance* inst);
void fri_commit_round(nlc_fri_state* s);
void fri_fold_round(nlc_fri_state* s);
void fri_final_poly_to_proof(nlc_fri_state* s);
size_t fri_final_poly_len(const nlc_fri_state* s);
void fri_queries(nlc_fri_state* s, size_t nq);
void fri_prove_openings(nlc_fri_state* s, const nlc_fri_instance* inst, nlc_challenger* ch_dev);
// plain view of the opaque FRI state for the C ABI layer
struct FriView {
    nlc_ctx* ctx;
    u64 *alpha, *beta, *pow, *resp, *xidx, *proof;
    size_t proof_words, final_off, pow_off, query_off, query_words, N;
    int round, cap_height, num_queries;
};
FriView fri_view(nlc_fri_state* s);
