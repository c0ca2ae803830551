/*
 * nlc_commit.h -- C ABI of libnlc_commit.so: the B200-native commitment backend behind plonky2's
 * PolynomialBatch / MerkleTree / fft helpers / FRI prover, as consumed by the unchanged starkyx and
 * plonky2x provers of near/near-light-client.
 *
 * Every entry point is `extern "C"`, takes plain pointers and sizes, returns 0 on success and a
 * non-zero code on failure (message via nlc_last_error).  There is NO CPU fallback: without a CUDA
 * device nlc_init fails and nothing else can be called.  plonky2's own contract at this boundary is
 * infallible-or-panic, so the Rust wrapper turns a non-zero code into panic!.
 *
 * "Replaces" cites the reference-side interface each entry point stands in for.  The plonky2 crate
 * itself (0.1.4 @ d2598bd, /root/reference/Cargo.lock:4864-4866) is not vendored in the reference
 * tree, so upstream paths are given by crate path and the in-tree CALL SITE by file:line.
 *
 * Conventions: field elements are uint64_t, canonical (< p = 2^64 - 2^32 + 1) on output; inputs may
 * be any representative.  A "column" is one polynomial: n = 2^log_n contiguous uint64_t.
 * Extension elements (D = 2) are two consecutive uint64_t [a0, a1].  A digest (HashOut) is 4 uint64_t.
 * Pointers are HOST pointers unless the name ends in _dev (device pointers on the context's GPU).
 * All entry points are thread-safe; calls on one context serialise on its stream.
 */
#ifndef NLC_COMMIT_H
#define NLC_COMMIT_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nlc_ctx nlc_ctx;     /* one per (process, GPU) */
typedef struct nlc_batch nlc_batch; /* a committed PolynomialBatch living in HBM */
typedef struct nlc_tree nlc_tree;   /* a stand-alone MerkleTree living in HBM */

/* ---- lifecycle ---------------------------------------------------------------------------- */
/* device < 0: the device named by the environment variable NLC_DEVICE if set, else the current CUDA device. */
int nlc_init(int device, nlc_ctx** out);
/* Every nlc_batch / nlc_tree / nlc_fri_state made on the context must have been freed before (the Rust wrapper's
 * handles hold an Arc of the context); a handle that outlives its context is dangling. */
void nlc_destroy(nlc_ctx* ctx);
/* Message of the CALLING THREAD's last failed call (threads sharing a context each see their own);
 * ctx may be NULL: error of this thread's last failed nlc_init.  Valid until the thread's next failing call. */
const char* nlc_last_error(nlc_ctx* ctx);
int nlc_device(nlc_ctx* ctx);
/* Launch all work of this context on an existing CUDA stream (a cudaStream_t cast to void*); NULL
 * restores the context's own stream.  Lets a harness time kernels with events on that stream. */
int nlc_set_stream(nlc_ctx* ctx, void* cuda_stream);
int nlc_synchronize(nlc_ctx* ctx);
/* STREAM AND LIFETIME CONTRACT of every *_dev entry point (nlc_commit_from_values_dev, nlc_commit_from_coeffs_dev,
 * nlc_commit_from_values_shard_dev, nlc_lde_columns[_scatter]_dev, nlc_commit_from_lde_rows_dev,
 * nlc_quotient_commit_dev, nlc_air_eval_dev): they only ENQUEUE work on the context's stream and return.  That stream is
 * created cudaStreamNonBlocking: it does NOT order against the legacy NULL stream or any other stream.  The caller must
 * (a) order the producer of the input buffer before the call -- nlc_wait_stream(ctx, producer) does it with one event
 * (NULL = the legacy default stream), or run the library on the producer's stream with nlc_set_stream -- and
 * (b) keep input and output device buffers alive and unmodified until nlc_synchronize(ctx) or a later synchronising
 * accessor (nlc_batch_cap, ...) has returned. */
int nlc_wait_stream(nlc_ctx* ctx, void* producer_stream);
/* Device buffers are cached per context by size (budget NLC_CACHE_MB, default 16384): out[0] = requests
 * served from the cache, out[1] = requests that went to the driver (cudaMallocAsync), out[2] = microseconds spent in
 * those driver calls, out[3] = bytes waiting in the cache now. */
int nlc_cache_stats(nlc_ctx* ctx, uint64_t out[4]);
/* Number of kernels this library has launched on ctx since creation (bench.py's gpu_launches). */
uint64_t nlc_launch_count(nlc_ctx* ctx);

/* plonky2's `timing: &mut TimingTree` argument (prover.rs:91-93): per-phase device time of the commits
 * issued since the last read, from CUDA events on the launching stream.  Phases follow upstream's scope
 * names: "IFFT", "FFT + blinding" (+ "transpose LDEs", fused), "build Merkle tree" split into leaf hashing
 * and the digest levels. */
typedef struct {
    double ifft_ms;
    double lde_ms;
    double leaf_hash_ms;
    double merkle_levels_ms;
    uint64_t commits;
} nlc_timing;
int nlc_timing_enable(nlc_ctx* ctx, int on);
int nlc_timing_read(nlc_ctx* ctx, nlc_timing* out); /* synchronises the stream, then resets */

/* ---- a1 / a2: PolynomialBatch::from_values / from_coeffs ----------------------------------- */
/* Replaces plonky2::fri::oracle::PolynomialBatch::from_values(values, rate_bits, blinding=false,
 * cap_height, timing, None); call sites vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:91-93 and
 * plonky2/stark/config.rs:92-94.  cols[j] = the n evaluations of polynomial j over the subgroup. */
int nlc_commit_from_values(nlc_ctx* ctx, const uint64_t* const* cols, size_t n_cols, int log_n, int rate_bits,
                           int cap_height, nlc_batch** out);
/* Replaces PolynomialBatch::from_coeffs(polynomials, rate_bits, false, cap_height, timing, None);
 * call site vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:171-178. */
int nlc_commit_from_coeffs(nlc_ctx* ctx, const uint64_t* const* cols, size_t n_cols, int log_n, int rate_bits,
                           int cap_height, nlc_batch** out);
/* from_values that ALSO fills the host images of the struct fields a patched plonky2 keeps, with the
 * device->host traffic hidden under the commit: coeffs_out_cols = `polynomials` (n_cols host vectors of n, what
 * proof.rs:307-310 reads), leaves_out = merkle_tree.leaves flattened (N x n_cols row-major), digests_out =
 * merkle_tree.digests, cap_out = merkle_tree.cap; any of them may be NULL.  The coefficients of a column chunk cross PCIe
 * as soon as its iFFT is done, while later chunks upload and transform; the leaf rows (whole only once every column is
 * extended) cross slab by slab under the leaf hashing and the digest levels; the digests + cap at the end.  Same results
 * as nlc_commit_from_values followed by nlc_batch_coeffs / _leaves / _digests / _cap (which pay the copies after the
 * commit). */
int nlc_commit_from_values_materialised(nlc_ctx* ctx, const uint64_t* const* cols, size_t n_cols, int log_n, int rate_bits,
                                        int cap_height, uint64_t* const* coeffs_out_cols, uint64_t* leaves_out,
                                        uint64_t* digests_out, uint64_t* cap_out, nlc_batch** out);
/* Same, input already resident in HBM as one n_cols x n matrix (column j at values_dev + j*n). */
int nlc_commit_from_values_dev(nlc_ctx* ctx, const uint64_t* values_dev, size_t n_cols, int log_n, int rate_bits,
                               int cap_height, nlc_batch** out);
int nlc_commit_from_coeffs_dev(nlc_ctx* ctx, const uint64_t* coeffs_dev, size_t n_cols, int log_n, int rate_bits,
                               int cap_height, nlc_batch** out);
/* "Next" row: commit straight from a row-major AirTrace (trace/mod.rs:19-24) instead of
 * the serial column split of trace/mod.rs:270-285: rows x width host matrix, columns [col0, col0+n_cols). */
int nlc_commit_from_trace_rows(nlc_ctx* ctx, const uint64_t* rows, size_t width, size_t col0, size_t n_cols,
                               int log_n, int rate_bits, int cap_height, nlc_batch** out);
/* SURVEY.md section 5 (checkpoint / resume): bring a PolynomialBatch that was built and serialised earlier back onto
 * the GPU -- plonky2x serialises circuit data incl. constants_sigmas_commitment (/root/reference/scripts/
 * build-circuit.sh:10-18), which is oracle 0 of every plonk prove_openings.  coeff_cols = `polynomials` (n_cols host
 * vectors of n coefficients).  The LDE and the tree are RECOMPUTED from them (bit-identical by construction, no 8x larger
 * upload); expect_cap (2^cap_height x 4) and expect_digests (2*(N - 2^cap_height) x 4), when non-NULL, are the
 * serialised merkle_tree.cap / .digests: a mismatch fails loudly (non-zero, message in nlc_last_error). */
int nlc_batch_import(nlc_ctx* ctx, const uint64_t* const* coeff_cols, size_t n_cols, int log_n, int rate_bits,
                     int cap_height, const uint64_t* expect_cap, const uint64_t* expect_digests, nlc_batch** out);
/* SURVEY.md 8(e).2 -- one GPU's share of a commit, split by LDE coset.  n_shards (a power of two, at most
 * 2^rate_bits and at most 2^cap_height) ranks each call this with their shard_idx: the batch then holds the
 * leaves [shard_idx, shard_idx + 1) * N / n_shards (whole cosets = whole cap subtrees), their digests, and
 * 2^cap_height / n_shards cap entries; concatenating the shards' caps / digests / leaves in shard order gives exactly
 * the single-GPU MerkleCap / digests / leaves.  The only exchange is the all-gather of the cap entries (32 bytes
 * each), done by the caller (NCCL all_gather in near-light-client_b200/plonky2.py). */
int nlc_commit_from_values_shard(nlc_ctx* ctx, const uint64_t* const* cols, size_t n_cols, int log_n, int rate_bits,
                                 int cap_height, size_t shard_idx, size_t n_shards, nlc_batch** out);
int nlc_commit_from_values_shard_dev(nlc_ctx* ctx, const uint64_t* values_dev, size_t n_cols, int log_n,
                                     int rate_bits, int cap_height, size_t shard_idx, size_t n_shards,
                                     nlc_batch** out);
/* The same split without replicating the iFFT or the upload (SURVEY.md 8e.2 "column-sharded iFFT + all-gather of the
 * coefficient matrix"): rank p uploads and inverse-transforms ONLY its columns [col0, col0 + n_cols) and the transform's
 * own stores put every coefficient into all n_peers coefficient matrices (peer_coeffs[q] = rank q's n_cols_total x n
 * buffer, nlc_peer_alloc / nlc_peer_open; this rank's own buffer included) -- the all-gather rides on the last NTT pass
 * over NVLink, no NCCL call.  After one barrier in stream order every rank holds the whole matrix and calls
 * nlc_commit_from_coeffs_shard_dev for its cosets; the batch reads coeffs_dev in place (it must outlive the batch). */
int nlc_ifft_columns_bcast_dev(nlc_ctx* ctx, const uint64_t* values_dev, size_t n_cols, int log_n, size_t col0,
                               size_t n_peers, uint64_t* const* peer_coeffs);
int nlc_commit_from_coeffs_shard_dev(nlc_ctx* ctx, uint64_t* coeffs_dev, size_t n_cols, int log_n, int rate_bits,
                                     int cap_height, size_t shard_idx, size_t n_shards, nlc_batch** out);
/* SURVEY.md 8(e).3 -- more GPUs than cosets (starky commits at rate_bits 1 on 4 or 8 GPUs): shard by COLUMN for the
 * transforms, exchange once, shard by leaf ROW for the tree.
 *   1. every rank, for the columns it owns: nlc_lde_columns_dev -- iFFT (is_values) and the 2^rate_bits coset NTTs of
 *      n_cols columns; lde_out_dev = n_cols x N column-major in leaf order (leaf k = LDE row bitrev_N(k)),
 *      coeffs_out_dev = n_cols x n or NULL.  Device pointers, asynchronous on the context's stream.
 *   2. all-to-all over NVLink (caller: NCCL all_to_all_single in near-light-client_b200/plonky2.py): rank g receives
 *      the leaves [g, g+1) * N / n_shards of every column -> n_cols_total x (N / n_shards) column-major.
 *   3. every rank: nlc_commit_from_lde_rows_dev over what it received -- leaf hashes, the digests of its cap
 *      subtrees and its 2^cap_height / n_shards cap entries, exactly as a coset shard (same concatenation rule); then
 *      the all-gather of the cap entries.  The batch does NOT copy lde_rows_dev: the buffer must outlive it.  It holds
 *      no coefficients (nlc_batch_coeffs / nlc_batch_eval fail on it; the column owners kept them in step 1).
 * n_shards: a power of two, at most 2^cap_height. */
int nlc_lde_columns_dev(nlc_ctx* ctx, const uint64_t* src_dev, size_t n_cols, int log_n, int rate_bits, int is_values,
                        uint64_t* coeffs_out_dev, uint64_t* lde_out_dev);
int nlc_commit_from_lde_rows_dev(nlc_ctx* ctx, uint64_t* lde_rows_dev, size_t n_cols, int log_n, int rate_bits,
                                 int cap_height, size_t shard_idx, size_t n_shards, nlc_batch** out);
/* Steps 1 and 2 fused: the last NTT pass stores each leaf where its row shard lives instead of into a local LDE.
 *   peer_rows == NULL: send_dev = [n_shards][n_cols][N / n_shards], the all-to-all's send buffer (no re-blocking pass);
 *   peer_rows != NULL: n_shards device pointers (host array), peer_rows[g] = rank g's receive buffer
 *     (n_cols_total x N / n_shards, opened with nlc_peer_open; this rank's own buffer for g = its rank): local column j
 *     is written to peer_rows[g] + (col0 + j) * (N / n_shards) over NVLink while the transform is still running --
 *     no send buffer, no NCCL call; the caller separates the transforms from the tree with one barrier.
 * n_shards <= 16. */
int nlc_lde_columns_scatter_dev(nlc_ctx* ctx, const uint64_t* src_dev, size_t n_cols, int log_n, int rate_bits,
                                int is_values, uint64_t* coeffs_out_dev, uint64_t* send_dev, size_t n_shards,
                                uint64_t* const* peer_rows, size_t col0);
/* Receive buffers that other processes on the node can write (cudaMalloc + CUDA IPC).  handle: 64 bytes to pass to
 * the peers (any transport); nlc_peer_open maps a peer's buffer into this process (peer access enabled lazily). */
int nlc_peer_alloc(nlc_ctx* ctx, size_t words, uint64_t** ptr_out, unsigned char handle_out[64]);
int nlc_peer_open(nlc_ctx* ctx, const unsigned char handle[64], uint64_t** ptr_out);
int nlc_peer_close(nlc_ctx* ctx, uint64_t* ptr);
int nlc_peer_free(nlc_ctx* ctx, uint64_t* ptr);
int nlc_batch_shard(const nlc_batch* b, size_t* shard_idx, size_t* n_shards);
void nlc_batch_free(nlc_batch* b);

/* Struct fields of PolynomialBatch / MerkleTree, materialised on demand. */
int nlc_batch_info(const nlc_batch* b, size_t* n_cols, int* log_n, int* rate_bits, int* cap_height);
/* merkle_tree.cap: 2^cap_height x 4 (prover.rs:95,180,210). */
int nlc_batch_cap(nlc_batch* b, uint64_t* out);
/* polynomials: column j written to out_cols[j] (n each), natural coefficient order (proof.rs:307-310). */
int nlc_batch_coeffs(nlc_batch* b, uint64_t* const* out_cols);
/* merkle_tree.leaves flattened: N x n_cols row-major, leaf k = LDE row bitrev_N(k).  The device keeps the LDE
 * column-major (the order the NTT produces); this call transposes it slab by slab on its way to the host. */
int nlc_batch_leaves(nlc_batch* b, uint64_t* out);
/* merkle_tree.digests: 2*(N - 2^cap_height) x 4 in plonky2's interleaved layout. */
int nlc_batch_digests(nlc_batch* b, uint64_t* out);
/* a6: get_lde_values(i, step) = leaves[bitrev_N(i*step)] for i in idx (prover.rs:289-294): m x n_cols. */
int nlc_batch_lde_rows(nlc_batch* b, const uint64_t* idx, size_t m, size_t step, uint64_t* out);
/* The same for the stripe i = i_start .. i_start + count - 1 in ONE launch and ONE D2H (count x n_cols): what a
 * host-side quotient loop (prover.rs:306-352 calls get_lde_values_packed(i, step) for every P::WIDTH points, twice)
 * uses while the AIR is still evaluated on the CPU; with nlc_air_eval_dev the rows never leave the GPU. */
int nlc_batch_lde_rows_range(nlc_batch* b, size_t i_start, size_t count, size_t step, uint64_t* out);
/* merkle_tree.get(leaf_index) and merkle_tree.prove(leaf_index): row (n_cols) and siblings
 * ((log N - cap_height) x 4, bottom-up). Either output may be NULL. */
int nlc_batch_open(nlc_batch* b, size_t leaf_index, uint64_t* row_out, uint64_t* siblings_out);
/* Device views (valid until nlc_batch_free): for callers that keep consuming on the GPU.
 * lde: n_cols x N column-major, lde[j*N + k] = column j of leaf k (leaf k = LDE row bitrev_N(k)). */
const uint64_t* nlc_batch_lde_dev(const nlc_batch* b);
const uint64_t* nlc_batch_coeffs_dev(const nlc_batch* b);
const uint64_t* nlc_batch_cap_dev(const nlc_batch* b);
/* StarkOpeningSet::new's eval_commitment (proof.rs:307-310): every polynomial of the batch evaluated
 * at the extension point z -> n_cols x 2. */
int nlc_batch_eval(nlc_batch* b, const uint64_t z[2], uint64_t* out);

/* ---- a3: fft helpers (plonky2_field::fft, polynomial::{PolynomialValues,PolynomialCoeffs}) -- */
/* In place on `n_cols` host columns stored back to back (data + j*n); natural order in and out.
 * Replace fft_with_options / ifft_with_options / coset_fft / coset_ifft; in-tree call sites
 * prover.rs:356 (coset_ifft(F::coset_shift())) and inside lde_onto_coset (prover.rs:280-284). */
int nlc_fft(nlc_ctx* ctx, uint64_t* data, size_t n_cols, int log_n);
int nlc_ifft(nlc_ctx* ctx, uint64_t* data, size_t n_cols, int log_n);
int nlc_coset_fft(nlc_ctx* ctx, uint64_t* data, size_t n_cols, int log_n, uint64_t shift);
int nlc_coset_ifft(nlc_ctx* ctx, uint64_t* data, size_t n_cols, int log_n, uint64_t shift);
/* PolynomialValues::lde_onto_coset(rate_bits): in = n_cols x n values, out = n_cols x (n << rate_bits). */
int nlc_lde_onto_coset(nlc_ctx* ctx, const uint64_t* in, uint64_t* out, size_t n_cols, int log_n, int rate_bits);

/* ---- "next" row SURVEY.md 8f.1: quotient_polys without the AIR evaluation -------------------- */
/* The tables vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:278-302 builds before its evaluation loop, over the
 * quotient domain of size = n << quotient_degree_bits (any output may be NULL):
 *   coset_x[i]        = g * w_size^i              (cyclic_subgroup_coset_known_order, prover.rs:298-302)
 *   lagrange_first/last = PolynomialValues::selector(n, 0 / n-1).lde_onto_coset(quotient_degree_bits) (prover.rs:278-283)
 *   zh_inv[k], k < 2^quotient_degree_bits: ZeroPolyOnCoset::new(log_n, quotient_degree_bits).eval_inverse(i) =
 *                       zh_inv[i mod 2^quotient_degree_bits]   (prover.rs:285, 336-337) */
int nlc_quotient_domain(nlc_ctx* ctx, int log_n, int quotient_degree_bits, uint64_t* coset_x, uint64_t* lagrange_first,
                        uint64_t* lagrange_last, uint64_t* zh_inv);
/* What prover.rs does with the constraint accumulators once the AIR has been evaluated (prover.rs:336-357 and
 * 158-178): evals = size x n_challenges row-major (one row per point of the quotient domain, natural order: the
 * `quotient_values` before `transpose`); if divide_by_zh the rows are first multiplied by 1/Z_H (pass the raw
 * accumulators), else they are taken as already divided.  Each column is coset_ifft(F::coset_shift())'d, checked
 * by trim_to_len(n * quotient_degree_factor) (error "Quotient has failed, the vanishing polynomial is not divisible
 * by Z_H" like the reference's expect), split into quotient_degree_factor chunks of n coefficients (polynomial j,
 * chunk c -> column j*factor + c) and committed as PolynomialBatch::from_coeffs(chunks, rate_bits, false,
 * cap_height, ..).  Errors if quotient_degree_bits > rate_bits (the assert of prover.rs:270-273). */
int nlc_quotient_commit(nlc_ctx* ctx, const uint64_t* evals, size_t n_challenges, int log_n, int quotient_degree_bits,
                        size_t quotient_degree_factor, int divide_by_zh, int rate_bits, int cap_height,
                        nlc_batch** out);
/* Same with the accumulators already in HBM (an on-device AIR evaluator, SURVEY.md 8f.3, feeds this). */
int nlc_quotient_commit_dev(nlc_ctx* ctx, const uint64_t* evals_dev, size_t n_challenges, int log_n,
                            int quotient_degree_bits, size_t quotient_degree_factor, int divide_by_zh, int rate_bits,
                            int cap_height, nlc_batch** out);

/* ---- "next" row SURVEY.md 8f.3: AIR constraint evaluation on the device -------------------------------------- */
/* Replaces the evaluation loop of starkyx's quotient_polys (vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:306-352):
 * per point of the quotient domain, local/next rows from every trace commitment (get_lde_values_packed,
 * prover.rs:289-294), `stark.air().eval(&mut parser)` (prover.rs:334) into a ConstraintConsumer
 * (plonky2/parser/consumer.rs:54-78).  The AIR is handed over as a straight-line BYTECODE that the caller obtains by
 * running its `air.eval` once against a RECORDING AirParser (air/parser.rs: Var = virtual register id; the way
 * RecursiveStarkParser, plonky2/parser/mod.rs:30-38, records a circuit): any AIR written against the AirParser trait --
 * Chip<L> constraint lists (chip/mod.rs:58-66), ArithmeticConstraint (chip/arithmetic/mod.rs:13-50), the field
 * instructions (chip/field/add.rs, mul.rs, ...) -- compiles unchanged.  One instruction = 4 x uint32 {op, dst, a, b}: */
/* Instruction word 0 = opcode (bits 0-7) | flags (bits 8-10).  The interpreter keeps the previous instruction's result
 * in an accumulator: A_LAST / B_LAST take operand a / b from it instead of from a register, NO_STORE says every use of
 * this result comes before the next result is produced, so it is never written to its register (dst is ignored).
 * Flags are an optimisation the recording compiler may or may not use; a program without them is equally valid. */
#define NLC_AIR_FLAG_A_LAST 1u
#define NLC_AIR_FLAG_B_LAST 2u
#define NLC_AIR_FLAG_NO_STORE 4u
#define NLC_AIR_LOCAL 0                  /* dst = local_slice()[a]      (a indexes the columns of all trace commitments, concatenated) */
#define NLC_AIR_NEXT 1                   /* dst = next_slice()[a] */
#define NLC_AIR_PUBLIC 2                 /* dst = public_slice()[a] */
#define NLC_AIR_GLOBAL 3                 /* dst = global_slice()[a] */
#define NLC_AIR_CHALLENGE 4              /* dst = challenge_slice()[a] */
#define NLC_AIR_CONST 5                  /* dst = constant(pool[a]) */
#define NLC_AIR_ADD 6                    /* dst = add(reg a, reg b) */
#define NLC_AIR_SUB 7
#define NLC_AIR_MUL 8
#define NLC_AIR_NEG 9                    /* dst = neg(reg a) */
#define NLC_AIR_MUL_CONST 10             /* dst = mul_const(reg a, pool[b]) */
#define NLC_AIR_ADD_CONST 11
#define NLC_AIR_SUB_CONST 12
#define NLC_AIR_CONSTRAINT 13            /* constraint(reg a) */
#define NLC_AIR_CONSTRAINT_TRANSITION 14 /* constraint_transition(reg a): * z_last = x - g^(n-1) */
#define NLC_AIR_CONSTRAINT_FIRST_ROW 15  /* constraint_first_row(reg a): * lagrange_basis_first */
#define NLC_AIR_CONSTRAINT_LAST_ROW 16   /* constraint_last_row(reg a):  * lagrange_basis_last */
#define NLC_AIR_MULADD_CONST 17          /* dst = reg a * pool[b] + accumulator (the previous result): sums of constant multiples */
/* dst = sum of `a` terms reg * constant: the terms follow in ceil(a / 2) words {reg, constant VALUE (< 2^16), reg,
 * constant}; a reg field with NLC_AIR_DOT_TERM_LAST set is the accumulator.  The 96-bit products are summed unreduced
 * (at most 2^12 terms), one reduction at the end: an MDS row, a limb recombination. */
#define NLC_AIR_DOT_CONST 18
#define NLC_AIR_DOT_TERM_LAST 0x80000000u
typedef struct {
    const uint32_t* code; /* n_ops x 4 */
    uint32_t n_ops;
    uint32_t n_regs;      /* virtual registers (a register is written before it is read; registers may be reused) */
    const uint64_t* pool; /* n_pool 64-bit field constants */
    uint32_t n_pool;
    uint32_t n_public;    /* lengths of the public / global / challenge slices the program may index */
    uint32_t n_global;
    uint32_t n_challenge;
} nlc_air_program;
/* constraint_accs of every point, NOT yet multiplied by 1/Z_H: evals_out_dev = size x n_alphas row-major in the natural
 * order of the quotient domain (size = n << quotient_degree_bits) -- exactly what nlc_quotient_commit_dev takes with
 * divide_by_zh = 1.  trace_data: the commitments whose LDE rows are concatenated into local/next vars (prover.rs:289-294);
 * all must share degree and rate_bits.  alphas: config.num_challenges values (prover.rs:262), at most 8.  Asynchronous
 * (see the *_dev contract above); nlc_air_eval is the same with the accumulators copied to the host. */
int nlc_air_eval_dev(nlc_ctx* ctx, const nlc_air_program* air, nlc_batch* const* trace_data, int n_trace_data,
                     const uint64_t* public_inputs, const uint64_t* global_values, const uint64_t* challenges,
                     const uint64_t* alphas, size_t n_alphas, int quotient_degree_bits, uint64_t* evals_out_dev);
int nlc_air_eval(nlc_ctx* ctx, const nlc_air_program* air, nlc_batch* const* trace_data, int n_trace_data,
                 const uint64_t* public_inputs, const uint64_t* global_values, const uint64_t* challenges,
                 const uint64_t* alphas, size_t n_alphas, int quotient_degree_bits, uint64_t* evals_out);
/* quotient_polys + the quotient commitment (prover.rs:144-178) without a host round trip: evaluate the AIR, scale by
 * 1/Z_H, per-alpha coset_ifft(F::coset_shift()), trim_to_len / chunks(degree), PolynomialBatch::from_coeffs(chunks,
 * rate_bits, false, cap_height).  quotient_degree_factor = stark.air().quotient_degree_factor(). */
int nlc_stark_quotient_commit(nlc_ctx* ctx, const nlc_air_program* air, nlc_batch* const* trace_data, int n_trace_data,
                              const uint64_t* public_inputs, const uint64_t* global_values, const uint64_t* challenges,
                              const uint64_t* alphas, size_t n_alphas, size_t quotient_degree_factor, int cap_height,
                              nlc_batch** out);

/* ---- the plonk-side quotient path (plonky2::plonk::{prover, vanishing_poly}) ----------------------
 * What `plonk::prover::prove` does between the wires commitment and the openings, with every LDE row read where it
 * already lives (call site in tree: /root/reference/nearx/src/test_utils.rs:62 `circuit.prove`, and every circuit the
 * operator builds, /root/reference/scripts/build-circuit.sh:10-18).  The circuit-dependent part -- the gates' constraint
 * polynomials -- arrives as an nlc_air_program: the caller evaluates every gate's FILTERED constraints ONCE against a
 * recording parser (the way `eval_unfiltered_circuit` records them for recursion), sums them per constraint index and
 * emits NLC_AIR_CONSTRAINT for the HIGHEST index first (near-light-client_b200/plonk.py: compile_gate_constraints).  In
 * that program LOCAL column j is constants_sigmas polynomial j for j < constants_sigmas width, then wire j - width;
 * PUBLIC i is public_inputs_hash[i]; NEXT and the row-filtered constraint ops are rejected. */
typedef struct {
    size_t num_constants;        /* polynomials of constants_sigmas before the sigmas: selectors, then gate constants */
    size_t num_routed_wires;     /* config.num_routed_wires; the sigmas are constants_sigmas[num_constants ..][..routed] */
    size_t num_challenges;       /* config.num_challenges: betas, gammas and alphas each hold this many (<= 8) */
    size_t num_partial_products; /* common_data.num_partial_products = ceil(routed / quotient_degree_factor) - 1 */
    int quotient_degree_factor;  /* common_data.quotient_degree_factor (the permutation argument's chunk size) */
    const uint64_t* k_is;        /* common_data.k_is: num_routed_wires coset shifts */
} nlc_plonk_circuit;
/* all_wires_permutation_partial_products + the prover's column order [Z per challenge | partial products per challenge]
 * + PolynomialBatch::from_values(zs_partial_products, rate_bits, false, cap_height): the wire and sigma values on the
 * subgroup are recovered from the two resident commitments (one forward NTT of their coefficients), the running
 * product Z(g x) = Z(x) * row product is a parallel prefix product over the rows. */
int nlc_plonk_zs_partial_products_commit(nlc_ctx* ctx, const nlc_plonk_circuit* circuit, nlc_batch* constants_sigmas,
                                         nlc_batch* wires, const uint64_t* betas, const uint64_t* gammas, int cap_height,
                                         nlc_batch** out);
/* eval_vanishing_poly_base_batch for every point of the quotient domain (size = n << log2_ceil(quotient_degree_factor)):
 * evals_out = size x num_challenges row-major, natural order, NOT yet multiplied by 1/Z_H (as nlc_air_eval). */
int nlc_plonk_vanishing_eval(nlc_ctx* ctx, const nlc_plonk_circuit* circuit, const nlc_air_program* gate_constraints,
                             nlc_batch* constants_sigmas, nlc_batch* wires, nlc_batch* zs_partial_products,
                             const uint64_t public_inputs_hash[4], const uint64_t* betas, const uint64_t* gammas,
                             const uint64_t* alphas, uint64_t* evals_out);
/* compute_quotient_polys + the quotient commitment of `prove`: the evaluation above, the 1/Z_H scaling, per-challenge
 * coset_ifft(F::coset_shift()), trim_to_len(quotient_degree) / chunks(degree), from_coeffs(chunks, rate_bits, false,
 * cap_height) -- num_challenges * quotient_degree_factor polynomials. */
int nlc_plonk_quotient_commit(nlc_ctx* ctx, const nlc_plonk_circuit* circuit, const nlc_air_program* gate_constraints,
                              nlc_batch* constants_sigmas, nlc_batch* wires, nlc_batch* zs_partial_products,
                              const uint64_t public_inputs_hash[4], const uint64_t* betas, const uint64_t* gammas,
                              const uint64_t* alphas, int cap_height, nlc_batch** out);

/* ---- a4: Poseidon (plonky2::hash::{poseidon,hashing}) -------------------------------------- */
/* PoseidonPermutation::permute on `count` independent 12-lane states (in place). */
int nlc_poseidon_permute(nlc_ctx* ctx, uint64_t* states, size_t count);
/* PoseidonHash::hash_no_pad / hash_or_noop of `count` inputs of `len` elements each -> count x 4. */
int nlc_hash_no_pad(nlc_ctx* ctx, const uint64_t* in, size_t count, size_t len, uint64_t* out);
int nlc_hash_or_noop(nlc_ctx* ctx, const uint64_t* in, size_t count, size_t len, uint64_t* out);
/* PoseidonHash::two_to_one on `count` pairs: left/right count x 4 -> out count x 4. */
int nlc_two_to_one(nlc_ctx* ctx, const uint64_t* left, const uint64_t* right, size_t count, uint64_t* out);

/* ---- a5: MerkleTree::new / prove (plonky2::hash::merkle_tree) ------------------------------ */
/* leaves: n_leaves x leaf_len row-major (host). */
int nlc_merkle_new(nlc_ctx* ctx, const uint64_t* leaves, size_t n_leaves, size_t leaf_len, int cap_height,
                   nlc_tree** out);
int nlc_tree_cap(nlc_tree* t, uint64_t* out);
int nlc_tree_digests(nlc_tree* t, uint64_t* out);
int nlc_tree_open(nlc_tree* t, size_t leaf_index, uint64_t* leaf_out, uint64_t* siblings_out);
void nlc_tree_free(nlc_tree* t);

/* ---- no allocation: Merkle cap with static device scratch ----------------------------------- */
/* No counterpart in the reference; added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * nlc_merkle_new + nlc_tree_cap (MerkleTree::new(leaves, cap_height).cap), but:
 *   - leaves_dev (n_leaves x leaf_len row-major) and cap_out_dev (2^cap_height x 4) are DEVICE pointers of the caller;
 *   - the digest layers and the leaf-unit counter live in a file-scope __device__ array of the library
 *     (csrc/static_scratch.cu: 4 MiB + 256 bytes, held by every GPU that loads the library until the process exits),
 *     so n_leaves <= NLC_STATIC_MAX_LEAVES; larger trees are rejected (non-zero, message in nlc_last_error);
 *   - the call makes no cudaMalloc / cudaFree / allocating library call and does not synchronise: the kernels are
 *     enqueued on the context's stream (nlc_set_stream), cap_out_dev is valid once that stream reaches the call's end;
 *   - all calls share the one array per GPU: calls that overlap on the same GPU (two streams, two contexts on one
 *     device) race; the array keeps the previous call's contents, the library clears what must start at zero. */
#define NLC_STATIC_MAX_LEAVES 65536
int nlc_merkle_cap_static_dev(nlc_ctx* ctx, const uint64_t* leaves_dev, size_t n_leaves, size_t leaf_len,
                              int cap_height, uint64_t* cap_out_dev);

/* ---- a9: Challenger (plonky2::iop::challenger) -------------------------------------------- */
/* The transcript state crosses the boundary as plain data, exactly the fields of
 * Challenger { sponge_state, input_buffer, output_buffer }.  The Rust Challenger keeps running on
 * the host between calls; prove_openings advances it on the device (no host round trips inside). */
typedef struct {
    uint64_t sponge_state[12];
    uint64_t input_buffer[8];
    uint32_t input_len;
    uint64_t output_buffer[8];
    uint32_t output_len;
} nlc_challenger;
void nlc_challenger_init(nlc_challenger* c);
/* Challenger::observe_elements / observe_cap / observe_extension_elements (prover.rs:67,94-100,181,193) */
int nlc_challenger_observe(nlc_ctx* ctx, nlc_challenger* c, const uint64_t* elems, size_t n);
/* Challenger::get_n_challenges (n = 2 gives get_extension_challenge) (prover.rs:99,183) */
int nlc_challenger_get(nlc_ctx* ctx, nlc_challenger* c, size_t n, uint64_t* out);

/* ---- a7 / a8 / a10: PolynomialBatch::prove_openings and the FRI prover --------------------- */
/* FriParams { config: FriConfig { rate_bits, cap_height, proof_of_work_bits, num_query_rounds, .. },
 * hiding = false, degree_bits, reduction_arity_bits } (plonky2::fri::{FriConfig,FriParams});
 * knobs set at vendor/starkyx/starkyx/src/plonky2/stark/config.rs:64-70. */
#define NLC_MAX_FRI_ROUNDS 32 /* degree_bits <= 32 (two-adicity) and every round removes >= 1 bit */
typedef struct {
    int degree_bits;
    int rate_bits;
    int cap_height;
    int proof_of_work_bits;
    int num_query_rounds;
    int n_reduction_rounds;
    int reduction_arity_bits[NLC_MAX_FRI_ROUNDS];
} nlc_fri_params;
/* FriConfig::fri_params(degree_bits, false) with FriReductionStrategy::ConstantArityBits(arity_bits,
 * final_poly_bits) (config.rs:68 uses (4, 5); config.rs:75-77).  Upstream's loop is unbounded; this one fails
 * (non-zero) rather than truncate if it would need more than NLC_MAX_FRI_ROUNDS rounds. */
int nlc_fri_params_constant_arity(nlc_fri_params* p, int degree_bits, int rate_bits, int cap_height,
                                  int proof_of_work_bits, int num_query_rounds, int arity_bits,
                                  int final_poly_bits);

/* FriInstanceInfo { oracles, batches: [FriBatchInfo { point, polynomials: [FriPolynomialInfo] }] }
 * (built at vendor/starkyx/starkyx/src/plonky2/stark/mod.rs:56-96). */
typedef struct {
    int n_batches;
    const uint64_t* points;        /* n_batches x 2 */
    const uint32_t* batch_offsets; /* n_batches + 1, into the two arrays below */
    const uint32_t* poly_oracle;   /* FriPolynomialInfo::oracle_index */
    const uint32_t* poly_index;    /* FriPolynomialInfo::polynomial_index */
} nlc_fri_instance;

/* Size in uint64_t words of the flat FriProof written by nlc_prove_openings:
 *   commit_phase_merkle_caps : n_rounds x 2^cap_height x 4
 *   final_poly               : (2^degree_bits >> sum(arity_bits)) x 2
 *   pow_witness              : 1
 *   query_round_proofs       : num_query_rounds x { per oracle: leaf row (n_cols), siblings ((logN-cap) x 4);
 *                                                   per round: evals (2 << arity), siblings ((logN_i - arity - cap) x 4) }
 */
size_t nlc_fri_proof_words(const nlc_fri_params* p, nlc_batch* const* oracles, int n_oracles);
/* Replaces PolynomialBatch::prove_openings(&instance, &oracles, &mut challenger, &fri_params, timing);
 * call site vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:200-206.  The proof-of-work witness is
 * the SMALLEST valid nonce (upstream: rayon find_any, any valid nonce). */
int nlc_prove_openings(nlc_ctx* ctx, const nlc_fri_instance* inst, nlc_batch* const* oracles, int n_oracles,
                       nlc_challenger* challenger, const nlc_fri_params* p, uint64_t* proof_out);

/* Finer-grained a8 steps for a Rust-driven commit phase (transcript stays in Rust between calls). */
typedef struct nlc_fri_state nlc_fri_state;
/* alpha-batched reduction + divide-by-linear + LDE + coset FFT over F_{p^2}: the state then holds
 * lde_polynomial_coeffs / lde_polynomial_values of fri_proof(). */
int nlc_fri_begin(nlc_ctx* ctx, const nlc_fri_instance* inst, nlc_batch* const* oracles, int n_oracles,
                  const uint64_t alpha[2], const nlc_fri_params* p, nlc_fri_state** out);
/* One fri_committed_trees iteration, first half: bit-reverse, chunk by arity, MerkleTree::new -> cap. */
int nlc_fri_commit_round(nlc_fri_state* s, uint64_t* cap_out);
/* Second half: fold coefficients by beta, shift ^= arity, coset FFT. */
int nlc_fri_fold_round(nlc_fri_state* s, const uint64_t beta[2]);
/* final_poly coefficients ((n >> sum arity) x 2). */
int nlc_fri_final_poly(nlc_fri_state* s, uint64_t* out);
/* fri_proof_of_work: grind on the challenger's duplex state; returns the smallest witness. */
int nlc_fri_pow(nlc_ctx* ctx, const nlc_challenger* c, int proof_of_work_bits, uint64_t* witness_out);
/* fri_prover_query_rounds for the given x_index values -> the query_round_proofs part of the flat proof. */
int nlc_fri_queries(nlc_fri_state* s, const uint64_t* x_indices, size_t n_queries, uint64_t* out);
void nlc_fri_state_free(nlc_fri_state* s);

#ifdef __cplusplus
}
#endif
#endif /* NLC_COMMIT_H */
