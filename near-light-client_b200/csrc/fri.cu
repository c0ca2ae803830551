// fri.cu -- PolynomialBatch::prove_openings and the FRI prover on the device.
//
// Replaces (plonky2 0.1.4 @ d2598bd; sources absent from /root/reference):
//   plonky2/src/fri/oracle.rs     PolynomialBatch::prove_openings
//   plonky2/src/util/reducing.rs  ReducingFactor::{reduce_polys_base, shift_poly}
//   field/src/polynomial/division.rs  PolynomialCoeffs::divide_by_linear
//   plonky2/src/fri/prover.rs     fri_proof, fri_committed_trees, fri_proof_of_work, fri_prover_query_rounds
//   plonky2/src/iop/challenger.rs Challenger (device-resident transcript, transcript.cuh)
// In-tree call site: /root/reference/vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:200-206.
//
// Extension-field polynomials live as two base-field planes (a0[], a1[]): plonky2's FFT over F_{p^2}
// uses base-field roots, so it is exactly two base NTTs.  The values of every FRI round are produced by
// the NTT directly in index-bit-reversed order with the two planes interleaved ([a0,a1] per element),
// which IS the leaf matrix of that round's Merkle tree (cur/arity rows of 2*arity words): no
// bit-reversal, chunking or flattening pass.  The whole of prove_openings is enqueued on one stream --
// transcript, folds, trees, proof of work and query gathers -- and the flat proof comes back in one copy.
#include <cstring>

#include "batch.cuh"
#include "transcript.cuh"

namespace {

using gl::Ext2;

inline Ext2 h_ext_mul(Ext2 x, Ext2 y) {
    auto add = [](u64 a, u64 b) { return gl::add(a, b); };
    return {add(gl::h_mul(x.a0, y.a0), gl::h_mul(7, gl::h_mul(x.a1, y.a1))),
            add(gl::h_mul(x.a0, y.a1), gl::h_mul(x.a1, y.a0))};
}
inline Ext2 h_ext_pow(Ext2 b, u64 e) {
    Ext2 r = {1, 0};
    while (e) {
        if (e & 1) r = h_ext_mul(r, b);
        b = h_ext_mul(b, b);
        e >>= 1;
    }
    return r;
}

// ---- transcript kernels (one warp) ----
__global__ void k_ch_observe(nlc_challenger* c, const u64* elems, size_t n) {
    transcript::observe(c, elems, n, threadIdx.x);
}
__global__ void k_ch_get(nlc_challenger* c, size_t n, u64* out) {
    for (size_t i = 0; i < n; i++) {
        u64 v = transcript::get(c, threadIdx.x);
        if (threadIdx.x == 0) out[i] = v;
    }
}

// ---- fri_proof_of_work: smallest nonce whose duplex response has >= pow_bits leading zeros ----
__global__ void __launch_bounds__(256) k_pow(const nlc_challenger* c, int pow_bits, unsigned long long* best) {
    u64 base_state[12];
#pragma unroll
    for (int i = 0; i < 12; i++) base_state[i] = c->sponge_state[i];
    const int pos = (int)c->input_len;
#pragma unroll
    for (int i = 0; i < 8; i++)
        if (i < pos) base_state[i] = c->input_buffer[i];
    const u64 stride = (u64)gridDim.x * blockDim.x;
    const u64 gtid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    for (u64 base = 0;; base += stride) {
        if (base > *(volatile unsigned long long*)best) break; // every smaller nonce has been claimed
        const u64 nonce = base + gtid;
        u64 s[12];
#pragma unroll
        for (int i = 0; i < 12; i++) s[i] = base_state[i];
#pragma unroll
        for (int i = 0; i < 8; i++)
            if (i == pos) s[i] = nonce;
        poseidon::permute_nc(s);
        const u64 resp = gl::canon(s[7]); // squeeze().last() == state[RATE - 1]
        if (pow_bits == 0 || (resp >> (64 - pow_bits)) == 0) atomicMin(best, (unsigned long long)nonce);
        if (base + stride < base) break;
    }
}

// out[k] = alpha^k, k < count (extension elements, interleaved)
__global__ void k_ext_pows(const u64* alpha, size_t count, u64* out) {
    const size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (k >= count) return;
    Ext2 a = {alpha[0], alpha[1]};
    Ext2 r = gl::ext_pow(a, k);
    out[2 * k] = r.a0;
    out[2 * k + 1] = r.a1;
}

// reduce_polys_base: comp[i] = sum_k alpha^k * poly_k[i].  A streaming pass over every committed coefficient
// (221 MB for a 422-column STARK): the polynomials are cut into slices (blockIdx.y) so that enough warps are in
// flight to cover HBM latency; slice sums are added by k_reduce_sum.
__global__ void __launch_bounds__(256)
k_reduce_polys(const u64* const* __restrict__ polys, const u64* __restrict__ apow, u32 K, u32 k_per, size_t n, u64* p0,
               u64* p1) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 k0 = blockIdx.y * k_per, k1 = k0 + k_per < K ? k0 + k_per : K;
    u64 acc0 = 0, acc1 = 0;
#pragma unroll 4
    for (u32 k = k0; k < k1; k++) { // the loads do not depend on the accumulators: several in flight per thread
        const u64 c = gl::canon(polys[k][i]);
        acc0 = gl::add(acc0, gl::mul(apow[2 * k], c));
        acc1 = gl::add(acc1, gl::mul(apow[2 * k + 1], c));
    }
    p0[blockIdx.y * n + i] = acc0;
    p1[blockIdx.y * n + i] = acc1;
}
__global__ void k_reduce_sum(const u64* p0, const u64* p1, u32 S, size_t n, u64* c0, u64* c1) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    u64 a0 = 0, a1 = 0;
    for (u32 s = 0; s < S; s++) {
        a0 = gl::add(a0, p0[s * n + i]);
        a1 = gl::add(a1, p1[s * n + i]);
    }
    c0[i] = a0;
    c1[i] = a1;
}

// Exclusive suffix Horner scan E[i] = sum_{j>i} x[j] rho^(j-i-1), three kernels + recursion.
// divide_by_linear(z) followed by the pushed zero is exactly E with x = coefficients, rho = z.
__global__ void k_scan_local(const u64* x0, const u64* x1, size_t chunks, int L, Ext2 rho, u64* s0, u64* s1) {
    const size_t c = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (c >= chunks) return;
    Ext2 acc = {0, 0};
    for (size_t j = (c + 1) * L; j-- > c * L;) {
        acc = gl::ext_mul(acc, rho);
        acc = gl::ext_add(acc, Ext2{x0[j], x1[j]});
    }
    s0[c] = acc.a0;
    s1[c] = acc.a1;
}
__global__ void k_scan_final(const u64* x0, const u64* x1, const u64* ec0, const u64* ec1, size_t chunks, int L,
                             Ext2 rho, u64* e0, u64* e1) {
    const size_t c = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (c >= chunks) return;
    Ext2 e = {ec0[c], ec1[c]};
    for (size_t i = (c + 1) * L; i-- > c * L;) {
        const Ext2 x = {x0[i], x1[i]};
        e0[i] = e.a0;
        e1[i] = e.a1;
        e = gl::ext_add(gl::ext_mul(e, rho), x);
    }
}
__global__ void k_scan_serial(const u64* x0, const u64* x1, size_t len, Ext2 rho, u64* e0, u64* e1) {
    if (blockIdx.x | threadIdx.x) return;
    Ext2 e = {0, 0};
    for (size_t i = len; i-- > 0;) {
        const Ext2 x = {x0[i], x1[i]};
        e0[i] = e.a0;
        e1[i] = e.a1;
        e = gl::ext_add(gl::ext_mul(e, rho), x);
    }
}
// shift_poly + add: f[i] = f[i] * shift + q[i]
__global__ void k_shift_add(u64* f0, u64* f1, const u64* q0, const u64* q1, const u64* shift, size_t n) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    Ext2 s = {shift[0], shift[1]};
    Ext2 f = gl::ext_add(gl::ext_mul(Ext2{f0[i], f1[i]}, s), Ext2{q0[i], q1[i]});
    f0[i] = f.a0;
    f1[i] = f.a1;
}
// coeffs'[i] = reduce_with_powers(coeffs[i*arity .. (i+1)*arity], beta)
__global__ void k_fold(const u64* c0, const u64* c1, size_t n_out, int arity_bits, const u64* beta, u64* o0, u64* o1) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n_out) return;
    const Ext2 b = {beta[0], beta[1]};
    const size_t arity = (size_t)1 << arity_bits;
    Ext2 acc = {0, 0};
    for (size_t k = arity; k-- > 0;) acc = gl::ext_add(gl::ext_mul(acc, b), Ext2{c0[i * arity + k], c1[i * arity + k]});
    o0[i] = acc.a0;
    o1[i] = acc.a1;
}
__global__ void k_interleave(const u64* c0, const u64* c1, size_t n, u64* out) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[2 * i] = c0[i];
    out[2 * i + 1] = c1[i];
}
__global__ void k_mask_indices(const u64* rands, size_t n, u64 mask, u64* out) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) out[i] = rands[i] & mask;
}

inline unsigned blocks_for(size_t n, unsigned t = 256) { return (unsigned)((n + t - 1) / t); }

// E = exclusive suffix Horner scan of x (planes, length len) with ratio rho.
void suffix_scan(nlc_ctx* ctx, const u64* x0, const u64* x1, size_t len, Ext2 rho, u64* e0, u64* e1) {
    const int L = 32;
    if (len <= 64 || len % L) { // one thread, one step per element: only for the top of the recursion
        NLC_LAUNCH(ctx, k_scan_serial, 1, 32, 0, x0, x1, len, rho, e0, e1);
        return;
    }
    const size_t chunks = len / L;
    u64* s = ctx->dmalloc(4 * chunks);
    u64 *s0 = s, *s1 = s + chunks, *ec0 = s + 2 * chunks, *ec1 = s + 3 * chunks;
    NLC_LAUNCH(ctx, k_scan_local, blocks_for(chunks, 128), 128, 0, x0, x1, chunks, L, rho, s0, s1);
    suffix_scan(ctx, s0, s1, chunks, h_ext_pow(rho, L), ec0, ec1);
    NLC_LAUNCH(ctx, k_scan_final, blocks_for(chunks, 128), 128, 0, x0, x1, ec0, ec1, chunks, L, rho, e0, e1);
    ctx->dfree(s);
}

} // namespace

void challenger_observe_dev(nlc_ctx* ctx, nlc_challenger* ch_dev, const u64* elems_dev, size_t n) {
    if (n) NLC_LAUNCH(ctx, k_ch_observe, 1, 32, 0, ch_dev, elems_dev, n);
}
void challenger_get_dev(nlc_ctx* ctx, nlc_challenger* ch_dev, size_t n, u64* out_dev) {
    if (n) NLC_LAUNCH(ctx, k_ch_get, 1, 32, 0, ch_dev, n, out_dev);
}
void fri_pow_dev(nlc_ctx* ctx, const nlc_challenger* ch_dev, int pow_bits, u64* witness_dev) {
    NLC_CUDA(cudaMemsetAsync(witness_dev, 0xFF, sizeof(u64), ctx->stream));
    // one wave of nonces ~ the expected search length 2^pow_bits: a bigger grid only hashes nonces nobody needs
    size_t ctas = pow_bits >= 8 && pow_bits < 40 ? (((size_t)1 << pow_bits) + 255) / 256 : 1;
    if (ctas > (size_t)ctx->sm_count * 3) ctas = (size_t)ctx->sm_count * 3;
    NLC_LAUNCH(ctx, k_pow, (unsigned)ctas, 256, 0, ch_dev, pow_bits, (unsigned long long*)witness_dev);
}

// ---------------------------------------------------------------------------------------------
struct nlc_fri_state {
    nlc_ctx* ctx = nullptr;
    nlc_fri_params p{};
    std::vector<nlc_batch*> oracles;
    int log_N = 0;
    size_t N = 0;
    int round = 0;     // rounds committed so far
    int folded = 0;    // rounds folded so far
    int cur_log = 0;   // log2 of the current codeword length
    u64 shift = gl::GENERATOR;
    u64* coeffs[2] = {nullptr, nullptr}; // ping-pong, two planes of stride N each
    int cur_buf = 0;
    std::vector<nlc_tree> trees;
    u64* small = nullptr; // [0..2) alpha, [2..4) beta, [4] pow witness, [5] pow response, [8..) query rands / indices
    u64* proof = nullptr; // flat proof on the device
    size_t proof_words = 0, final_off = 0, pow_off = 0, query_off = 0, query_words = 0;
    u64* alpha_dev() { return small; }
    u64* beta_dev() { return small + 2; }
    u64* pow_dev() { return small + 4; }
    u64* resp_dev() { return small + 5; }
    u64* rands_dev() { return small + 8; }
    u64* xidx_dev() { return small + 8 + p.num_query_rounds; }
    u64* plane(int buf, int comp) { return coeffs[buf] + (size_t)comp * N; }
};

static int total_arity(const nlc_fri_params& p) {
    int t = 0;
    for (int i = 0; i < p.n_reduction_rounds; i++) t += p.reduction_arity_bits[i];
    return t;
}

size_t fri_proof_words_impl(const nlc_fri_params* p, nlc_batch* const* oracles, int n_oracles, size_t* final_off,
                            size_t* pow_off, size_t* query_off, size_t* query_words) {
    const int log_N = p->degree_bits + p->rate_bits;
    size_t w = (size_t)p->n_reduction_rounds * ((size_t)4 << p->cap_height);
    if (final_off) *final_off = w;
    w += (size_t)2 << (p->degree_bits - total_arity(*p));
    if (pow_off) *pow_off = w;
    w += 1;
    if (query_off) *query_off = w;
    size_t q = 0;
    for (int o = 0; o < n_oracles; o++) q += oracles[o]->W + (size_t)4 * (log_N - p->cap_height);
    int lg = log_N;
    for (int i = 0; i < p->n_reduction_rounds; i++) {
        const int ab = p->reduction_arity_bits[i];
        q += ((size_t)2 << ab) + (size_t)4 * (lg - ab - p->cap_height);
        lg -= ab;
    }
    if (query_words) *query_words = q;
    return w + q * (size_t)p->num_query_rounds;
}

void fri_state_release(nlc_fri_state* s) {
    if (!s) return;
    nlc_ctx* ctx = s->ctx;
    ctx->dfree(s->coeffs[0]);
    ctx->dfree(s->coeffs[1]);
    for (auto& t : s->trees) {
        ctx->dfree(t.leaves);
        ctx->dfree(t.digests);
    }
    ctx->dfree(s->small);
    ctx->dfree(s->proof);
    delete s;
}

// Values of the current coefficient vector on shift*<w_cur>, bit-reversed and interleaved = round leaves.
static void fri_values_to_leaves(nlc_fri_state* s, u64* leaves_out) {
    nlc_ctx* ctx = s->ctx;
    u64* scratch = nullptr;
    const size_t sw = ntt_scratch_words(2, s->cur_log);
    if (sw) scratch = ctx->dmalloc(sw);
    NttJob job;
    job.src = s->plane(s->cur_buf, 0);
    job.src_stride = s->N;
    job.n_cols = 2;
    job.log_n = s->cur_log;
    job.pre_shift = s->shift;
    job.rows = true;
    job.dst = leaves_out;
    job.row_width = 2;
    job.scratch = scratch;
    ntt_run(ctx, job);
    ctx->dfree(scratch);
}

nlc_fri_state* fri_state_new(nlc_ctx* ctx, const nlc_fri_instance* inst, nlc_batch* const* oracles, int n_oracles,
                             const nlc_fri_params* p) {
    NLC_CHECK(n_oracles > 0 && inst && p, "prove_openings: no oracles");
    const int log_n = p->degree_bits, log_N = log_n + p->rate_bits;
    NLC_CHECK(log_n >= 0 && log_N <= 32, "degree_bits + rate_bits out of range");
    NLC_CHECK(p->n_reduction_rounds >= 0 && p->n_reduction_rounds <= NLC_MAX_FRI_ROUNDS, "too many reduction rounds");
    NLC_CHECK(total_arity(*p) <= log_n, "reduction arities exceed the degree");
    NLC_CHECK(p->num_query_rounds >= 0 && p->proof_of_work_bits >= 0 && p->proof_of_work_bits < 64, "bad FriConfig");
    for (int o = 0; o < n_oracles; o++) {
        NLC_CHECK(oracles[o] && oracles[o]->ctx == ctx, "oracle belongs to another context");
        NLC_CHECK(oracles[o]->shard_log == 0, "prove_openings needs whole batches, not coset shards");
        NLC_CHECK(oracles[o]->coeffs != nullptr, "prove_openings needs the coefficients: not a leaf-row shard");
        NLC_CHECK(oracles[o]->log_n == log_n && oracles[o]->rate_bits == p->rate_bits &&
                      oracles[o]->cap_height == p->cap_height,
                  "oracle degree / rate_bits / cap_height do not match the FriParams");
    }
    {
        int lg = log_N;
        for (int i = 0; i < p->n_reduction_rounds; i++) {
            lg -= p->reduction_arity_bits[i];
            NLC_CHECK(p->reduction_arity_bits[i] >= 1 && p->reduction_arity_bits[i] <= 8, "arity bits out of range");
            NLC_CHECK(lg >= p->cap_height, "cap_height should be at most log2(leaves.len())");
        }
    }
    for (int b = 0; b < inst->n_batches; b++)
        for (u32 k = inst->batch_offsets[b]; k < inst->batch_offsets[b + 1]; k++) {
            NLC_CHECK((int)inst->poly_oracle[k] < n_oracles, "FriPolynomialInfo: oracle_index out of range");
            NLC_CHECK(inst->poly_index[k] < oracles[inst->poly_oracle[k]]->W,
                      "FriPolynomialInfo: polynomial_index out of range");
        }
    nlc_fri_state* s = new nlc_fri_state();
    s->ctx = ctx;
    s->p = *p;
    s->oracles.assign(oracles, oracles + n_oracles);
    s->log_N = log_N;
    s->N = (size_t)1 << log_N;
    s->cur_log = log_N;
    try {
        s->coeffs[0] = ctx->dmalloc(2 * s->N);
        s->coeffs[1] = ctx->dmalloc(2 * s->N);
        s->small = ctx->dmalloc(8 + 2 * (size_t)p->num_query_rounds + 8);
        s->proof_words =
            fri_proof_words_impl(p, oracles, n_oracles, &s->final_off, &s->pow_off, &s->query_off, &s->query_words);
        s->proof = ctx->dmalloc(s->proof_words);
    } catch (...) {
        fri_state_release(s);
        throw;
    }
    return s;
}

// alpha must already be in s->alpha_dev().  Computes lde_final_poly (coeffs) and lde_final_values.
void fri_begin(nlc_fri_state* s, const nlc_fri_instance* inst) {
    nlc_ctx* ctx = s->ctx;
    const size_t n = (size_t)1 << s->p.degree_bits, N = s->N;
    NLC_CUDA(cudaMemsetAsync(s->coeffs[0], 0, 2 * N * sizeof(u64), ctx->stream));
    u32 Kmax = 0;
    for (int b = 0; b < inst->n_batches; b++) Kmax = std::max(Kmax, inst->batch_offsets[b + 1] - inst->batch_offsets[b]);
    u64* apow = ctx->dmalloc(2 * ((size_t)Kmax + 1));
    NLC_LAUNCH(ctx, k_ext_pows, blocks_for((size_t)Kmax + 1), 256, 0, s->alpha_dev(), (size_t)Kmax + 1, apow);
    u64* tmp = ctx->dmalloc(4 * n); // comp planes + quotient planes
    u64 *c0 = tmp, *c1 = tmp + n, *q0 = tmp + 2 * n, *q1 = tmp + 3 * n;
    const u64** ptrs_dev = (const u64**)ctx->dmalloc((size_t)Kmax + 1);
    for (int b = 0; b < inst->n_batches; b++) {
        const u32 lo = inst->batch_offsets[b], hi = inst->batch_offsets[b + 1], K = hi - lo;
        std::vector<const u64*> ptrs(K);
        for (u32 k = 0; k < K; k++)
            ptrs[k] = s->oracles[inst->poly_oracle[lo + k]]->coeffs + (size_t)inst->poly_index[lo + k] * n;
        // the pointer table is consumed by the kernel launched right after; a pageable source makes this copy
        // complete (staged) before the call returns, so the vector may go out of scope
        if (K) NLC_CUDA(cudaMemcpyAsync(ptrs_dev, ptrs.data(), K * sizeof(u64*), cudaMemcpyHostToDevice, ctx->stream));
        const u32 S = K >= 64 ? std::min<u32>(16, K / 32) : 1, k_per = (K + S - 1) / S;
        if (S == 1) {
            NLC_LAUNCH(ctx, k_reduce_polys, dim3(blocks_for(n), 1), 256, 0, ptrs_dev, apow, K, K ? K : 1, n, c0, c1);
        } else {
            u64* part = ctx->dmalloc(2 * (size_t)S * n);
            NLC_LAUNCH(ctx, k_reduce_polys, dim3(blocks_for(n), S), 256, 0, ptrs_dev, apow, K, k_per, n, part,
                       part + (size_t)S * n);
            NLC_LAUNCH(ctx, k_reduce_sum, blocks_for(n), 256, 0, part, part + (size_t)S * n, S, n, c0, c1);
            ctx->dfree(part);
        }
        const Ext2 z = {gl::canon(inst->points[2 * b]), gl::canon(inst->points[2 * b + 1])};
        suffix_scan(ctx, c0, c1, n, z, q0, q1);
        NLC_LAUNCH(ctx, k_shift_add, blocks_for(n), 256, 0, s->plane(0, 0), s->plane(0, 1), q0, q1, apow + 2 * (size_t)K, n);
    }
    ctx->dfree((void*)ptrs_dev);
    ctx->dfree(tmp);
    ctx->dfree(apow);
    s->cur_buf = 0;
    s->cur_log = s->log_N;
    s->shift = gl::GENERATOR;
    s->round = s->folded = 0;
}

// fri_committed_trees, first half of an iteration: tree over the current values; cap lands in the proof.
void fri_commit_round(nlc_fri_state* s) {
    nlc_ctx* ctx = s->ctx;
    NLC_CHECK(s->round < s->p.n_reduction_rounds && s->round == s->folded, "fri_commit_round: out of order");
    const int ab = s->p.reduction_arity_bits[s->round];
    const size_t cur = (size_t)1 << s->cur_log;
    nlc_tree t;
    t.ctx = ctx;
    t.n_leaves = cur >> ab;
    t.leaf_len = (size_t)2 << ab;
    t.cap_height = s->p.cap_height;
    t.leaves = ctx->dmalloc(2 * cur);
    t.digests = ctx->dmalloc(8 * (t.n_leaves - ((size_t)1 << t.cap_height)));
    t.cap = s->proof + (size_t)s->round * ((size_t)4 << t.cap_height);
    s->trees.push_back(t);
    fri_values_to_leaves(s, t.leaves);
    merkle_build(ctx, t.leaves, t.n_leaves, t.leaf_len, 0, t.cap_height, t.digests, t.cap);
    s->round++;
}

// second half: fold by beta (in s->beta_dev()), shift <- shift^arity.
void fri_fold_round(nlc_fri_state* s) {
    nlc_ctx* ctx = s->ctx;
    NLC_CHECK(s->folded < s->round, "fri_fold_round: commit the round first");
    const int ab = s->p.reduction_arity_bits[s->folded];
    const size_t nxt = (size_t)1 << (s->cur_log - ab);
    const int nb = s->cur_buf ^ 1;
    NLC_LAUNCH(ctx, k_fold, blocks_for(nxt), 256, 0, s->plane(s->cur_buf, 0), s->plane(s->cur_buf, 1), nxt, ab,
               s->beta_dev(), s->plane(nb, 0), s->plane(nb, 1));
    s->cur_buf = nb;
    s->cur_log -= ab;
    s->shift = gl::h_pow(s->shift, (u64)1 << ab);
    s->folded++;
}

size_t fri_final_poly_len(const nlc_fri_state* s) { return (size_t)1 << (s->p.degree_bits - total_arity(s->p)); }

void fri_final_poly_to_proof(nlc_fri_state* s) {
    NLC_CHECK(s->folded == s->p.n_reduction_rounds, "final_poly: commit phase not finished");
    const size_t len = fri_final_poly_len(s);
    NLC_LAUNCH(s->ctx, k_interleave, blocks_for(len), 256, 0, s->plane(s->cur_buf, 0), s->plane(s->cur_buf, 1), len,
               s->proof + s->final_off);
}

// fri_prover_query_rounds for indices already in s->xidx_dev().
void fri_queries(nlc_fri_state* s, size_t nq) {
    nlc_ctx* ctx = s->ctx;
    NLC_CHECK(s->folded == s->p.n_reduction_rounds, "queries: commit phase not finished");
    u64* out = s->proof + s->query_off;
    size_t off = 0;
    for (nlc_batch* b : s->oracles) {
        merkle_open_many(ctx, b->lde, b->digests, s->N, b->W, s->N, b->cap_height, s->xidx_dev(), 0, nq, out,
                         s->query_words, off);
        off += b->W + (size_t)4 * (s->log_N - b->cap_height);
    }
    int sh = 0, lg = s->log_N;
    for (int r = 0; r < s->p.n_reduction_rounds; r++) {
        const int ab = s->p.reduction_arity_bits[r];
        sh += ab;
        const nlc_tree& t = s->trees[r];
        merkle_open_many(ctx, t.leaves, t.digests, t.n_leaves, t.leaf_len, 0, t.cap_height, s->xidx_dev(), sh, nq, out,
                         s->query_words, off);
        off += t.leaf_len + (size_t)4 * (lg - ab - t.cap_height);
        lg -= ab;
    }
}

// PolynomialBatch::prove_openings, everything on the stream; ch_dev is the device copy of the challenger.
void fri_prove_openings(nlc_fri_state* s, const nlc_fri_instance* inst, nlc_challenger* ch_dev) {
    nlc_ctx* ctx = s->ctx;
    const nlc_fri_params& p = s->p;
    challenger_get_dev(ctx, ch_dev, 2, s->alpha_dev());
    fri_begin(s, inst);
    const size_t cap_words = (size_t)4 << p.cap_height;
    for (int r = 0; r < p.n_reduction_rounds; r++) {
        fri_commit_round(s);
        challenger_observe_dev(ctx, ch_dev, s->proof + (size_t)r * cap_words, cap_words);
        challenger_get_dev(ctx, ch_dev, 2, s->beta_dev());
        fri_fold_round(s);
    }
    fri_final_poly_to_proof(s);
    challenger_observe_dev(ctx, ch_dev, s->proof + s->final_off, 2 * fri_final_poly_len(s));
    fri_pow_dev(ctx, ch_dev, p.proof_of_work_bits, s->pow_dev());
    challenger_observe_dev(ctx, ch_dev, s->pow_dev(), 1);
    challenger_get_dev(ctx, ch_dev, 1, s->resp_dev());
    NLC_CUDA(cudaMemcpyAsync(s->proof + s->pow_off, s->pow_dev(), sizeof(u64), cudaMemcpyDeviceToDevice, ctx->stream));
    const size_t nq = (size_t)p.num_query_rounds;
    if (nq) {
        challenger_get_dev(ctx, ch_dev, nq, s->rands_dev());
        NLC_LAUNCH(ctx, k_mask_indices, blocks_for(nq), 256, 0, s->rands_dev(), nq, (u64)(s->N - 1), s->xidx_dev());
        fri_queries(s, nq);
    }
}

FriView fri_view(nlc_fri_state* s) {
    FriView v;
    v.ctx = s->ctx;
    v.alpha = s->alpha_dev();
    v.beta = s->beta_dev();
    v.pow = s->pow_dev();
    v.resp = s->resp_dev();
    v.xidx = s->xidx_dev();
    v.proof = s->proof;
    v.proof_words = s->proof_words;
    v.final_off = s->final_off;
    v.pow_off = s->pow_off;
    v.query_off = s->query_off;
    v.query_words = s->query_words;
    v.N = s->N;
    v.round = s->round;
    v.cap_height = s->p.cap_height;
    v.num_queries = s->p.num_query_rounds;
    return v;
}
